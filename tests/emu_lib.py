"""TEST INFRASTRUCTURE ONLY: builds and loads tests/emu/_build/librtcuda_emu.so — the product's CUDA sources compiled
for the host against the SIMT emulation shim (tests/emu/shim/cuda_runtime.h) — so that the kernels' logic can be
compared with the oracle in a container without a GPU.  Never used by myraytracer_b200 itself."""
import ctypes as C
import os
import subprocess

from myraytracer_b200 import capi

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "emu")
EMU_PATH = os.path.join(EMU_DIR, "_build", "librtcuda_emu.so")
_emu = None


def load():
    global _emu
    if _emu is None:
        r = subprocess.run(["make", "-j8"], cwd=EMU_DIR, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("emulation build failed:\n" + r.stdout[-4000:])
        _emu = capi.declare(C.CDLL(EMU_PATH))
        _emu.rtemu_kernel_launches.restype = C.c_uint64
        _emu.rtemu_set_schedule_seed.argtypes = [C.c_uint64]
        _emu.rtemu_set_schedule_seed.restype = None
        _emu.rtemu_schedule_shuffles.restype = C.c_uint64
    return _emu


class use_emulator:
    """Context manager: inside it, myraytracer_b200.capi.load() hands out the emulation library."""

    def __enter__(self):
        self._saved = capi._lib
        capi._lib = load()
        # RTEMU_SCHEDULE_SEED=n runs a whole test session under a shuffled block / fiber order (emu_runtime.cpp)
        capi._lib.rtemu_set_schedule_seed(int(os.environ.get("RTEMU_SCHEDULE_SEED", "0")))
        return capi._lib

    def __exit__(self, *a):
        capi._lib.rtemu_set_schedule_seed(0)
        capi._lib = self._saved
