"""Mirror of reference ``bigvgan/env.py:8-11`` (``AttrDict``)."""


class AttrDict(dict):
    """dict whose keys are also attributes (reference env.py:8-11)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.__dict__ = self
