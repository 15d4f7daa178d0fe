// Host-side helpers shared between the .cu translation units of libdtds_b200.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace dtds {
// Number of SMs of the current device (cached per device).
int sm_count();
// 2-D bf16 tensor map with 128-byte swizzle; dim0 is the contiguous dimension.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1,
                      uint64_t ld_elems, uint32_t box0, uint32_t box1);
}  // namespace dtds
