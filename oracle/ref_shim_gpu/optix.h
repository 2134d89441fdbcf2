// Stand-in for the un-vendored OWL / OptiX header of the same name.
// Test infrastructure only: lets the reference device TU compile with nvcc for sm_100a (oracle/_ref/libref_gpu_*.so).
#pragma once
#include "owl_shim_gpu.h"
