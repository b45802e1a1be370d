// ORACLE shim (test infrastructure): stands in for NVIDIA's <optix.h>, which is not in this image.
// The reference's host-compilable headers only need the traversable-handle typedef.
#pragma once
typedef unsigned long long OptixTraversableHandle;
