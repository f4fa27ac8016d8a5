/* Umbrella header of the CRTCuda SwiftPM target (same pattern as Sources/CPly/include/CPly.h:5-6). */
#ifndef CRTCUDA_H
#define CRTCUDA_H
#include "rtcuda.h"
#include "rtcuda_host.h"
#endif
