// src/poly_gpu.cpp — drop-in replacement for the reference's src/poly.cpp (15 lines: from_blob, serialize,
// deserialize) against the unmodified kzg.h.  poly::from_blob is the most expensive CPU step per blob once the
// commitment itself runs on the GPU (≈1 s for 4096 chunks with NTL); for the consecutive nodes
// every kzg::blob has, the interpolation runs on the device (kzg_gpu_interpolate_range).  A blob built by hand
// from arbitrary x-coordinates keeps the reference's own host routine (polyfit, util.cpp:180-184).
#include <kzg.h>

#include "gpu_side.h"
#include "util.h"

kzg::poly kzg::poly::from_blob(kzg::blob blob) {
  int64_t first_x = 0;
  std::vector<uint8_t> ys;
  if (kzg_b200::consecutive_nodes(blob.get_data(), first_x, ys)) {
    ZZ_pX I;
    kzg_b200::interpolate_consecutive(first_x, ys, &I, nullptr);
    return kzg::poly(I);
  }
  return kzg::poly(polyfit(blob.get_data()));
}

std::vector<uint8_t> kzg::poly::serialize() { return serialize_ZZ_pX(data); }

kzg::poly kzg::poly::deserialize(const std::vector<uint8_t>& bytes) { return kzg::poly(deserialize_ZZ_pX(bytes)); }
