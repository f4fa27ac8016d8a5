// shim.cpp — the one compiled source of the CRTCuda SwiftPM target.
// SwiftPM cannot compile .cu files, so this target only re-exports the C ABI of the prebuilt
// librtcuda.so (built by `make -C myraytracer_b200/csrc`, nvcc -gencode arch=compute_100a,code=sm_100a)
// and makes sure the library is linked.  Same packaging pattern as the reference's CPly target
// (Package.swift:21-36: path, sources, publicHeadersPath "include", module.modulemap with an umbrella header).
#include "CRTCuda.h"

extern "C" int32_t crtcuda_abi_version(void) { return rt_abi_version(); }
