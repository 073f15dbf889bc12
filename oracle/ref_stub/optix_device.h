/* Empty stand-in for the OptiX SDK header (absent from this image).  The reference's
 * include/Global/DeviceFunctions.cuh includes <optix_device.h> but its host-callable maths
 * uses nothing from it.  Used only by oracle/Makefile's `_ref` target. */
#pragma once
