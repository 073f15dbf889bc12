/* Stand-in for the reference's include/Global/HostFunctions.cuh.  TEST INFRASTRUCTURE ONLY (oracle/_ref).
 * The real header hard-includes SDL2, Vulkan and OptiX host headers (HostFunctions.cuh:20-27,47-48), all
 * absent from this image.  include/Util/ColorRamp.cuh needs from it only the standard library headers, the
 * CUDA vector types and SDL_Log (one error message, ColorRamp.cuh:148).  oracle/Makefile puts
 * oracle/ref_stub before the reference's include directory, so <Global/HostFunctions.cuh> resolves here
 * while every other <Global/...> header still comes from the reference checkout. */
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

static inline void SDL_Log(const char *, ...) {}
