#!/bin/bash
# Install the UNMODIFIED reference vocoder into the git-ignored baseline/_ref/ (it still travels to the GPU box with the
# gpurun snapshot; /root/reference itself does not exist there):
#   * the BigVGAN package indextts/s2mel/modules/bigvgan/ (pure Python/torch: the CPU arm of `bench.py --impl reference`
#     and the same-GPU stock-PyTorch baseline), copied byte for byte;
#   * its fused CUDA extension K1 (alias_free_activation/cuda/anti_alias_activation{.cpp,_cuda.cu}), built HERE for sm_100
#     with the reference's own flag set (load.py:36-56, the sm_70/sm_80 -gencode pairs replaced by sm_100) into
#     baseline/_ref/k1/anti_alias_activation_cuda.so, so that tools/bench_act1d.py can time it beside bvg_activation1d_fwd
#     on the B200 ("the bar to beat on the same box").
# Nothing under index-tts_b200/ imports baseline/_ref; only bench.py's reference arm, tools/ and tests/ do.
# A pip install of /root/reference is not used: the project pins dependencies that are absent here and the
# vocoder package needs nothing but torch.
set -uo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
DST="${HERE}/_ref"
SRC="${REF}/indextts/s2mel/modules/bigvgan"
if [ ! -d "${SRC}" ]; then echo "install_ref: ${SRC} not found (GPU box: uses the prebuilt baseline/_ref)"; exit 0; fi
PKG="${DST}/indextts/s2mel/modules/bigvgan"
mkdir -p "${DST}/indextts/s2mel/modules"
rm -rf "${PKG}"
cp -r "${SRC}" "${PKG}"
rm -rf "${PKG}/alias_free_activation/cuda/build" "${PKG}/__pycache__"
find "${PKG}" -name "__pycache__" -type d -exec rm -rf {} + 2>/dev/null
# namespace-package parents: empty on purpose (the reference's indextts/__init__.py pulls in the whole TTS stack)
echo "installed the reference BigVGAN package into ${PKG}"

K1="${DST}/k1"
mkdir -p "${K1}"
if [ ! -f "${K1}/anti_alias_activation_cuda.so" ] || [ "${SRC}/alias_free_activation/cuda/anti_alias_activation_cuda.cu" -nt "${K1}/anti_alias_activation_cuda.so" ]; then
  TORCH_CUDA_ARCH_LIST="" python - "$PKG" "$K1" <<'PY' || echo "install_ref: K1 build failed (bench_act1d.py will report it as unavailable)"
import sys, pathlib
from torch.utils import cpp_extension
pkg, out = pathlib.Path(sys.argv[1]), sys.argv[2]
src = pkg / "alias_free_activation" / "cuda"
cpp_extension.load(
    name="anti_alias_activation_cuda",
    sources=[str(src / "anti_alias_activation.cpp"), str(src / "anti_alias_activation_cuda.cu")],
    build_directory=out,
    extra_cflags=["-O3"],
    # the reference's flags (load.py:36-56) with the arch pair set to the B200
    extra_cuda_cflags=["-O3", "-gencode", "arch=compute_100,code=sm_100", "--use_fast_math",
                       "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__", "--expt-relaxed-constexpr",
                       "--expt-extended-lambda"],
    verbose=False, is_python_module=False)
print("built K1 for sm_100 in", out)
PY
fi
ls -la "${K1}" 2>/dev/null | head
exit 0
