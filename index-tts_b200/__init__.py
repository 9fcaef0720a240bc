"""b200-bigvgan: B200-native BigVGAN-v2 vocoder forward behind the reference's module surface.

Import as ``index_tts_b200`` (see ../index_tts_b200/__init__.py); mirrors
``indextts/s2mel/modules/bigvgan/`` of the reference: ``bigvgan``, ``activations``, ``env``, ``utils``,
``alias_free_activation.{torch,cuda}``.
"""
