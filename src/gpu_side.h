// Shared between the drop-in translation units (trusted_setup_gpu.cpp, poly_gpu.cpp): a process-wide device
// context for the polynomial-only entry points (no SRS needed), and the consecutive-node test.
#ifndef KZG_B200_GPU_SIDE_H
#define KZG_B200_GPU_SIDE_H
#include <kzg.h>

#include <cstdint>
#include <vector>

#include "../include/kzg_gpu.h"

namespace kzg_b200 {

// lazily created context of this build's curve on device 0; throws std::runtime_error without a GPU
kzg_gpu_ctx* poly_ctx();

// If the blob's x-coordinates are first_x, first_x + 1, ... (what blob::from_bytes / from_string produce,
// src/blob.cpp:20-48) with 0 <= first_x < 2^62, pack the y values as CURVE_ORDER_BYTES little-endian records and
// return true.
bool consecutive_nodes(const std::vector<std::pair<NTL::ZZ_p, NTL::ZZ_p>>& points, int64_t& first_x,
                       std::vector<uint8_t>& ys);

// (I, Z) through consecutive nodes on the device; Z may be nullptr
void interpolate_consecutive(int64_t first_x, const std::vector<uint8_t>& ys, NTL::ZZ_pX* I, NTL::ZZ_pX* Z);

}  // namespace kzg_b200
#endif
