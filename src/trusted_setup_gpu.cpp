// src/trusted_setup_gpu.cpp — drop-in replacement for the reference's src/trusted_setup.cpp.
//
// Compiled against the reference's UNMODIFIED headers (src/kzg.h, src/util.h, config/curve_*/kzg_config.h)
// and linked with the reference's unmodified blob/poly/commit/proof/util translation units, so every caller
// of kzg.h (benchmark/benchmark.cpp, testing/testing.cpp, demo/shared/kzg-cli.cpp) builds and runs unchanged.
// What changes is below the private methods (reference file:line in brackets):
//   polyeval_G1  [src/trusted_setup.cpp:149-174]   -> kzg_gpu_msm_g1      (Pippenger MSM on the B200)
//   polyeval_G2  [src/trusted_setup.cpp:176-201]   -> kzg_gpu_msm_g2
//   create_proof [src/trusted_setup.cpp:214-228]   -> kzg_gpu_prove       (quotient + MSM chained on device;
//                                                     (P - I)/Z == P div Z, so neither I nor the evaluations are needed)
//   trusted_setup(int) [src/trusted_setup.cpp:21-74, 123-135] -> kzg_gpu_setup_powers (fixed-base batch)
// What does not change: argument checks and exception types/messages, the SRS file format and the two pairings
// (PAIR_ate / PAIR_fexp, src/trusted_setup.cpp:242-253), which stay MIRACL host code by design.  verify_proof's
// interpolation (linear_roots_and_polyfit, util.cpp:172-178) runs on the device for consecutive nodes
// (kzg_gpu_interpolate_range) and keeps the reference's host routine for hand-built blobs with arbitrary x.
// There is no CPU fallback: a failing C-ABI call (no GPU, out of memory) throws std::runtime_error.
//
// kzg.h gives trusted_setup no spare member and no destructor, so the device context of a setup lives in
// a side table keyed by the object's address and validated by a fingerprint of its SRS.
#include <kzg.h>

#include <algorithm>
#include <array>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "gpu_side.h"
#include "util.h"
#include "../include/kzg_gpu.h"

int kzg::CURVE_ORDER_BYTES;

namespace {

constexpr int M = MODBYTES_CURVE;
constexpr int G1_BYTES = 2 * M;   // x | y
constexpr int G2_BYTES = 4 * M;   // x.b | x.a | y.b | y.a
constexpr int CURVE_ID = (M == 20) ? KZG_GPU_BN158 : (M == 32) ? KZG_GPU_BN254 : KZG_GPU_BLS12381;

[[noreturn]] void fail(const kzg_gpu_ctx* ctx, int rc, const char* what) {
  std::string msg = std::string("kzg gpu: ") + what + ": " + kzg_gpu_strerror(rc);
  if (ctx && *kzg_gpu_last_error(ctx)) msg += std::string(" [") + kzg_gpu_last_error(ctx) + "]";
  throw std::runtime_error(msg);
}

// affine big-endian coordinates of a point, i.e. the octet body after the 0x04 tag
void g1_body(ECP& p, uint8_t* out) {
  char buf[G1_BYTES + 1];
  octet o = {0, (int)sizeof(buf), buf};
  ECP_toOctet(&o, &p, false);          // normalises p in place; infinity leaves zeros
  if (ECP_isinf(&p)) memset(out, 0, G1_BYTES); else memcpy(out, buf + 1, G1_BYTES);
}
void g2_body(ECP2& p, uint8_t* out) {
  char buf[G2_BYTES + 1];
  octet o = {0, (int)sizeof(buf), buf};
  ECP2_toOctet(&o, &p, false);
  if (ECP2_isinf(&p)) memset(out, 0, G2_BYTES); else memcpy(out, buf + 1, G2_BYTES);
}
ECP g1_from_body(const uint8_t* xy, int is_inf) {
  ECP p;
  if (is_inf) { ECP_inf(&p); return p; }
  char buf[G1_BYTES + 1];
  buf[0] = 0x04;
  memcpy(buf + 1, xy, G1_BYTES);
  octet o = {G1_BYTES + 1, (int)sizeof(buf), buf};
  if (!ECP_fromOctet(&p, &o)) throw std::runtime_error("kzg gpu: device returned a point that is not on the curve");
  return p;
}
ECP2 g2_from_body(const uint8_t* xy, int is_inf) {
  ECP2 p;
  if (is_inf) { ECP2_inf(&p); return p; }
  char buf[G2_BYTES + 1];
  buf[0] = 0x04;
  memcpy(buf + 1, xy, G2_BYTES);
  octet o = {G2_BYTES + 1, (int)sizeof(buf), buf};
  if (!ECP2_fromOctet(&p, &o)) throw std::runtime_error("kzg gpu: device returned a G2 point that is not on the curve");
  return p;
}

// G2 point from an octet body the device has just produced in kzg_gpu_setup_powers, WITHOUT the on-curve test:
// ECP2_fromOctet -> ECP2_set -> ECP2_rhs divides the curve constant by the twist element on every call
// (FP2_div_ip -> FP_inv, ecp2.cpp:301-318, 268-298), ~20 us per point = most of trusted_setup(4096) once the
// scalar multiplications are on the GPU.  The result is the exact representation ECP2_set leaves (x, y, z = 1);
// the caller still runs the checked path on a few indices, and the reference checks every point when it loads
// an exported file.  All-zero coordinates = infinity (s = 0 only).
ECP2 g2_from_device_body(const uint8_t* xy) {
  ECP2 p;
  bool zero = true;
  for (int i = 0; i < G2_BYTES; i++) zero = zero && xy[i] == 0;
  if (zero) { ECP2_inf(&p); return p; }
  FP2_fromBytes(&p.x, (char*)xy);
  FP2_fromBytes(&p.y, (char*)xy + G2_BYTES / 2);
  FP2_one(&p.z);
  return p;
}

// coefficients of P as flat little-endian CURVE_ORDER_BYTES records (what BIG_from_ZZ starts from, util.cpp:31-43)
// BytesFromZZ only reads the coefficient's magnitude (no ZZ_p modulus context involved), so large polynomials are
// packed by a few host threads: at 0.2-0.4 us per coefficient the single-threaded loop (1.2 ms for 4097
// coefficients) would otherwise cost more than the 1.0 ms device MSM it feeds (SURVEY.md §7.4 "host marshalling").
std::vector<uint8_t> pack_coeffs(const ZZ_pX& P) {
  const long n = deg(P) + 1;
  const int SB = kzg::CURVE_ORDER_BYTES;
  std::vector<uint8_t> raw((size_t)n * SB);
  auto pack = [&](long lo, long hi) {
    for (long i = lo; i < hi; i++) BytesFromZZ(raw.data() + (size_t)i * SB, rep(P[i]), SB);
  };
  unsigned hw = std::thread::hardware_concurrency();
  const long nt = n < 2048 ? 1 : (long)std::min<unsigned>(hw ? hw : 4, 8);
  if (nt == 1) { pack(0, n); return raw; }
  std::vector<std::thread> pool;
  for (long t = 1; t < nt; t++) pool.emplace_back(pack, n * t / nt, n * (t + 1) / nt);
  pack(0, n / nt);
  for (auto& th : pool) th.join();
  return raw;
}

// ---- side table: trusted_setup object -> devices with its SRS resident
//
// kzg.h gives trusted_setup no spare member and no destructor, so the device side of a setup is found through two
// maps under one mutex:
//   by content   fingerprint of ALL its G1 and G2 bytes -> the resident copy (strong reference, least recently used
//                of more than KZG_GPU_CACHE = 8 dropped): copies of a trusted_setup, and setups loaded twice from one
//                file, share ONE resident SRS;
//   by address   object address -> weak reference + a cheap stamp (size, first/last G1, second/last G2), so the
//                common call does not re-serialise the SRS.
// Callers receive a shared_ptr: a context that is evicted or replaced while another thread is inside a call on it
// is destroyed only when that call returns.  Lookup, upload and insertion happen under the mutex, so two threads
// making the first use of one setup upload it once.  Points are serialised from COPIES (ECP_toOctet normalises
// its argument in place; the reference's readers never write to _G1/_G2 either).
struct Side {
  kzg_gpu_group* group = nullptr;
  size_t n = 0;
  ~Side() { if (group) kzg_gpu_group_destroy(group); }
  kzg_gpu_ctx* ctx() const { return kzg_gpu_group_ctx(group, 0); }
};
using SidePtr = std::shared_ptr<Side>;
using Digest = std::array<uint64_t, 2>;

struct ByAddress {
  std::weak_ptr<Side> side;
  size_t n = 0;
  std::vector<uint8_t> stamp;
};
struct ByContent {
  SidePtr side;
  uint64_t age = 0;
};
std::mutex g_mu;
std::map<const void*, ByAddress> g_by_address;
std::map<Digest, ByContent> g_by_content;
struct ByRaw {
  std::weak_ptr<Side> side;
  Digest content;
};
std::map<Digest, ByRaw> g_by_raw;  // in-memory fingerprint (raw_digest_of) -> resident copy: a shortcut to g_by_content
uint64_t g_clock = 0;

size_t cache_capacity() {
  const char* e = getenv("KZG_GPU_CACHE");
  const long v = e ? atol(e) : 0;
  return v >= 1 ? (size_t)v : 8;
}
// how many devices a setup of n points is made resident on: all of them when a single commitment is large enough for
// a split by point range to pay (kzg_gpu_group_msm_g1), one otherwise; KZG_GPU_DEVICES overrides
int devices_for(size_t n) {
  const int visible = kzg_gpu_device_count();
  if (visible <= 0) return 0;
  const char* e = getenv("KZG_GPU_DEVICES");
  if (e && atoi(e) >= 1) return std::min(atoi(e), visible);
  // One device below 2^20 points: a commitment over fewer points takes under ~4 ms, and every further device costs
  // a process 0.3-1 s of context creation (measured on 2 GPUs with the reference's one-operation-per-process demo
  // CLI at 65536 points: the whole drop-in test file 150 s -> 373 s).  From 2^20 points on a single create_commit
  // is split by point range over all visible devices (kzg_gpu_group_msm_g1: 3.6x on 8 GPUs at 2^20, 6.4x at 2^22).
  return n >= ((size_t)1 << 20) ? visible : 1;
}
// 128-bit fingerprint of a byte range: 64-bit words through two multiply-xorshift lanes.  Long inputs are cut into
// pieces hashed side by side on the host cores and the piece digests are folded in order (a 10.4 M-point SRS is 3 GB:
// a byte-at-a-time serial loop took 4.5 s of the constructor).
Digest digest_bytes(const uint8_t* p, size_t len) {
  Digest d = {0xcbf29ce484222325ull ^ len, 0x84222325cbf29ce4ull + len * 0x9e3779b97f4a7c15ull};
  auto mix = [&](uint64_t w) {
    d[0] = (d[0] ^ w) * 0x100000001b3ull;
    d[0] ^= d[0] >> 29;
    d[1] = (d[1] + w) * 0x9e3779b97f4a7c15ull;
    d[1] ^= d[1] >> 32;
  };
  size_t i = 0;
  for (; i + 8 <= len; i += 8) {
    uint64_t w;
    memcpy(&w, p + i, 8);
    mix(w);
  }
  uint64_t tail = 0;
  if (i < len) memcpy(&tail, p + i, len - i);
  mix(tail ^ ((uint64_t)(len - i) << 56));
  return d;
}
template <class PieceFn>  // piece(k, lo, hi) -> Digest of items [lo, hi)
Digest digest_in_pieces(size_t n_items, size_t min_items_per_piece, PieceFn piece) {
  unsigned hw = std::thread::hardware_concurrency();
  const size_t np = std::max<size_t>(1, std::min<size_t>(std::min<size_t>(hw ? hw : 4, 16), n_items / std::max<size_t>(1, min_items_per_piece)));
  std::vector<Digest> part(np);
  std::vector<std::thread> pool;
  for (size_t k = 1; k < np; k++) pool.emplace_back([&, k] { part[k] = piece(n_items * k / np, n_items * (k + 1) / np); });
  part[0] = piece(0, n_items / np);
  for (auto& th : pool) th.join();
  return digest_bytes(reinterpret_cast<const uint8_t*>(part.data()), part.size() * sizeof(Digest));
}
Digest digest_of(const std::vector<uint8_t>& b1, const std::vector<uint8_t>& b2) {
  auto of = [](const std::vector<uint8_t>& v) {
    return digest_in_pieces(v.size(), (size_t)1 << 20, [&](size_t lo, size_t hi) { return digest_bytes(v.data() + lo, hi - lo); });
  };
  const Digest a = of(b1), b = of(b2);
  const uint64_t both[6] = {a[0], a[1], b[0], b[1], (uint64_t)b1.size(), (uint64_t)b2.size()};
  return digest_bytes(reinterpret_cast<const uint8_t*>(both), sizeof both);
}
// Fingerprint of the points AS THEY LIE IN MEMORY (limbs and excess words of every coordinate, no padding): copies of
// a trusted_setup and the object a constructor just filled have the same one, and computing it needs no field
// arithmetic.  It is only ever used to FIND the resident copy of an equal object sooner (serialising 10.4 M points for
// the content fingerprint took 15 s in front of the first create_commit of a copied setup); two equal setups whose
// limbs differ (MIRACL's residues are not unique) simply miss here and are matched by content.
Digest raw_digest_of(const std::vector<ECP>& g1, const std::vector<ECP2>& g2) {
  auto fp = [](const FP& f, uint64_t* w) {
    memcpy(w, f.g, sizeof f.g);
    w[sizeof f.g / 8] = (uint64_t)(uint32_t)f.XES;
  };
  constexpr size_t FW = sizeof(((FP*)nullptr)->g) / 8 + 1;
  const Digest a = digest_in_pieces(g1.size(), 4096, [&](size_t lo, size_t hi) {
    std::vector<uint64_t> buf(3 * FW * 256);
    Digest d = {hi - lo, lo};
    for (size_t i = lo; i < hi;) {
      const size_t m = std::min<size_t>(256, hi - i);
      for (size_t j = 0; j < m; j++) {
        const ECP& p = g1[i + j];
        fp(p.x, &buf[(3 * j) * FW]); fp(p.y, &buf[(3 * j + 1) * FW]); fp(p.z, &buf[(3 * j + 2) * FW]);
      }
      const Digest e = digest_bytes(reinterpret_cast<const uint8_t*>(buf.data()), m * 3 * FW * 8);
      d[0] = (d[0] ^ e[0]) * 0x100000001b3ull;
      d[1] = (d[1] + e[1]) * 0x9e3779b97f4a7c15ull;
      i += m;
    }
    return d;
  });
  const Digest b = digest_in_pieces(g2.size(), 4096, [&](size_t lo, size_t hi) {
    std::vector<uint64_t> buf(6 * FW * 256);
    Digest d = {hi - lo, lo};
    for (size_t i = lo; i < hi;) {
      const size_t m = std::min<size_t>(256, hi - i);
      for (size_t j = 0; j < m; j++) {
        const ECP2& q = g2[i + j];
        uint64_t* o = &buf[6 * j * FW];
        fp(q.x.a, o); fp(q.x.b, o + FW); fp(q.y.a, o + 2 * FW); fp(q.y.b, o + 3 * FW); fp(q.z.a, o + 4 * FW); fp(q.z.b, o + 5 * FW);
      }
      const Digest e = digest_bytes(reinterpret_cast<const uint8_t*>(buf.data()), m * 6 * FW * 8);
      d[0] = (d[0] ^ e[0]) * 0x100000001b3ull;
      d[1] = (d[1] + e[1]) * 0x9e3779b97f4a7c15ull;
      i += m;
    }
    return d;
  });
  const uint64_t both[6] = {a[0], a[1], b[0], b[1], (uint64_t)g1.size(), (uint64_t)g2.size()};
  return digest_bytes(reinterpret_cast<const uint8_t*>(both), sizeof both);
}
std::vector<uint8_t> stamp_of(const std::vector<ECP>& g1, const std::vector<ECP2>& g2) {
  std::vector<uint8_t> s(2 * G1_BYTES + 2 * G2_BYTES, 0);
  ECP a = g1.front(), b = g1.back();  // copies: g1_body normalises in place
  g1_body(a, s.data());
  g1_body(b, s.data() + G1_BYTES);
  if (!g2.empty()) {
    ECP2 c = g2[g2.size() > 1 ? 1 : 0], d = g2.back();
    g2_body(c, s.data() + 2 * G1_BYTES);
    g2_body(d, s.data() + 2 * G1_BYTES + G2_BYTES);
  }
  return s;
}
SidePtr new_side(size_t n) {
  const int ndev = devices_for(n);
  SidePtr side = std::make_shared<Side>();
  int rc = ndev > 0 ? kzg_gpu_group_create(&side->group, CURVE_ID, nullptr, ndev) : KZG_GPU_ERR_CUDA;
  if (rc) fail(nullptr, rc, "kzg_gpu_group_create (a CUDA device is required; there is no CPU fallback)");
  side->n = n;
  // KZG_GPU_DIRECT_TABLE_GB=<GiB>: a small setup (the 4096-element blobs of src/blob.cpp) may also keep the direct-lookup
  // table of csrc/msm_direct.cuh — 32 GiB for 4097 points on BLS12-381 — which takes a degree-4096 create_commit from
  // 0.49 to 0.31 ms.  Off unless asked for: the table is per resident setup.
  if (const char* e = getenv("KZG_GPU_DIRECT_TABLE_GB"))
    if (atoi(e) > 0 && n <= 16384) kzg_gpu_group_set_option(side->group, "direct_table_gb", atoi(e));
  return side;
}
void remember_locked(const void* self, const SidePtr& side, const Digest& digest, std::vector<uint8_t> stamp) {
  ByContent& c = g_by_content[digest];
  c.side = side;
  c.age = ++g_clock;
  g_by_address[self] = ByAddress{side, side->n, std::move(stamp)};
  const size_t cap = cache_capacity();
  while (g_by_content.size() > cap) {  // drop the least recently used; in-flight callers keep theirs alive
    auto oldest = g_by_content.begin();
    for (auto it = g_by_content.begin(); it != g_by_content.end(); ++it)
      if (it->second.age < oldest->second.age) oldest = it;
    g_by_content.erase(oldest);
  }
  if (g_by_address.size() > 256)
    for (auto it = g_by_address.begin(); it != g_by_address.end();)
      it = it->second.side.expired() ? g_by_address.erase(it) : std::next(it);
  for (auto it = g_by_raw.begin(); it != g_by_raw.end();)
    it = it->second.side.expired() ? g_by_raw.erase(it) : std::next(it);
}
// device side of `self`, uploading the SRS the first time (or again if the object at this address changed)
SidePtr device_of(const void* self, const std::vector<ECP>& g1, const std::vector<ECP2>& g2) {
  if (g1.empty()) throw std::runtime_error("kzg gpu: empty trusted setup");
  std::lock_guard<std::mutex> lock(g_mu);
  std::vector<uint8_t> stamp = stamp_of(g1, g2);
  auto it = g_by_address.find(self);
  if (it != g_by_address.end() && it->second.n == g1.size() && it->second.stamp == stamp) {
    if (SidePtr side = it->second.side.lock()) {
      for (auto& kv : g_by_content)
        if (kv.second.side == side) kv.second.age = ++g_clock;
      return side;
    }
  }
  const size_t n = g1.size();
  const Digest raw = raw_digest_of(g1, g2);
  {
    auto rh = g_by_raw.find(raw);
    if (rh != g_by_raw.end())
      if (SidePtr side = rh->second.side.lock())
        if (side->n == n) {  // a copy of (or the very object behind) a resident setup: nothing to serialise
          remember_locked(self, side, rh->second.content, std::move(stamp));
          return side;
        }
  }
  std::vector<uint8_t> b1(n * (size_t)G1_BYTES), b2(g2.size() == n ? n * (size_t)G2_BYTES : 0);
  {
    auto pack = [&](size_t lo, size_t hi) {
      for (size_t i = lo; i < hi; i++) {
        ECP p = g1[i];
        g1_body(p, b1.data() + i * G1_BYTES);
        if (!b2.empty()) {
          ECP2 q = g2[i];
          g2_body(q, b2.data() + i * G2_BYTES);
        }
      }
    };
    unsigned hw = std::thread::hardware_concurrency();
    const size_t nt = n < 4096 ? 1 : std::min<size_t>(hw ? hw : 4, 16);
    std::vector<std::thread> pool;
    for (size_t t = 1; t < nt; t++) pool.emplace_back(pack, n * t / nt, n * (t + 1) / nt);
    pack(0, n / nt);
    for (auto& th : pool) th.join();
  }
  const Digest digest = digest_of(b1, b2);
  auto hit = g_by_content.find(digest);
  if (hit != g_by_content.end() && hit->second.side->n == n) {  // a copy of a setup that is already resident
    SidePtr side = hit->second.side;
    remember_locked(self, side, digest, std::move(stamp));
    g_by_raw[raw] = ByRaw{side, digest};
    return side;
  }
  SidePtr side = new_side(n);
  int rc = kzg_gpu_group_srs_load(side->group, n, b1.data(), b2.empty() ? nullptr : b2.data());
  if (rc) throw std::runtime_error(std::string("kzg gpu: srs_load failed: ") + kzg_gpu_group_last_error(side->group));
  remember_locked(self, side, digest, std::move(stamp));
  g_by_raw[raw] = ByRaw{side, digest};
  return side;
}

void wipe(void* p, size_t len) {  // not elidable
  volatile unsigned char* v = static_cast<volatile unsigned char*>(p);
  for (size_t i = 0; i < len; i++) v[i] = 0;
}

}  // namespace

// ---- shared with poly_gpu.cpp (gpu_side.h)
kzg_gpu_ctx* kzg_b200::poly_ctx() {
  static std::mutex mu;
  static kzg_gpu_ctx* ctx = nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!ctx) {
    int rc = kzg_gpu_ctx_create(&ctx, CURVE_ID, 0);
    if (rc) fail(nullptr, rc, "kzg_gpu_ctx_create (a CUDA device is required; there is no CPU fallback)");
  }
  return ctx;
}

bool kzg_b200::consecutive_nodes(const std::vector<std::pair<NTL::ZZ_p, NTL::ZZ_p>>& points, int64_t& first_x,
                                 std::vector<uint8_t>& ys) {
  if (points.empty()) return false;
  const ZZ& x0 = rep(points[0].first);
  if (NumBits(x0) > 62) return false;
  first_x = (int64_t)conv<long>(x0);
  const int SB = kzg::CURVE_ORDER_BYTES;
  ys.assign(points.size() * (size_t)SB, 0);
  for (size_t i = 0; i < points.size(); i++) {
    const ZZ& xi = rep(points[i].first);
    if (NumBits(xi) > 62 || (int64_t)conv<long>(xi) != first_x + (int64_t)i) return false;
    BytesFromZZ(ys.data() + i * SB, rep(points[i].second), SB);
  }
  return true;
}

void kzg_b200::interpolate_consecutive(int64_t first_x, const std::vector<uint8_t>& ys, NTL::ZZ_pX* I, NTL::ZZ_pX* Z) {
  const int SB = kzg::CURVE_ORDER_BYTES;
  const size_t k = ys.size() / SB;
  std::vector<uint8_t> ci(I ? k * SB : 0), cz(Z ? (k + 1) * SB : 0);
  kzg_gpu_ctx* ctx = poly_ctx();
  int rc = kzg_gpu_interpolate_range(ctx, ys.data(), first_x, k, I ? ci.data() : nullptr, Z ? cz.data() : nullptr);
  if (rc) fail(ctx, rc, "interpolate_range");
  auto unpack = [&](const std::vector<uint8_t>& raw, size_t n, NTL::ZZ_pX& out) {
    out.SetLength((long)n);
    ZZ z;
    for (size_t i = 0; i < n; i++) {
      ZZFromBytes(z, raw.data() + i * SB, SB);
      conv(out[(long)i], z);
    }
    out.normalize();
  };
  if (I) unpack(ci, k, *I);
  if (Z) unpack(cz, k + 1, *Z);
}

// ------------------------------------------------------------------------------------------------
void kzg::init() {  // same contract as the reference: NTL modulus = group order, CURVE_ORDER_BYTES = its byte length
  ZZ order = ZZ_from_BIG(CURVE_Order);
  ZZ_p::init(order);
  kzg::CURVE_ORDER_BYTES = NumBytes(order);
}

// trusted_setup(int): the secret is drawn exactly as the reference does (generate_random_BIG, util.cpp:62-76);
// the n powers and the 2n fixed-base multiplications run on the device, which keeps the SRS resident.
kzg::trusted_setup::trusted_setup(int num_coeff) {
  if (num_coeff < 2) throw invalid_argument("num_coeff must be at least 2");
  if (kzg::CURVE_ORDER_BYTES == 0) kzg::init();
  const int SB = kzg::CURVE_ORDER_BYTES;
  const size_t n = (size_t)num_coeff;
  std::vector<uint8_t> b1(n * G1_BYTES), b2(n * G2_BYTES);
  SidePtr side = new_side(n);
  int rc;
  {
    // The secret goes BIG -> big-endian bytes -> SB little-endian bytes without an NTL temporary, and both copies are
    // wiped before anything else happens; the device scrubs its copy and the powers before setup_powers returns.
    BIG secret;
    generate_random_BIG(secret);
    char be[MODBYTES_CURVE];
    BIG_toBytes(be, secret);
    uint8_t s_le[MODBYTES_CURVE];
    for (int i = 0; i < SB; i++) s_le[i] = (uint8_t)be[MODBYTES_CURVE - 1 - i];
    rc = kzg_gpu_group_setup_powers(side->group, s_le, n, b1.data(), b2.data());
    wipe(secret, sizeof(secret));
    wipe(be, sizeof(be));
    wipe(s_le, sizeof(s_le));
  }
  if (rc) throw std::runtime_error(std::string("kzg gpu: setup_powers failed: ") + kzg_gpu_group_last_error(side->group));
  _G1.resize(n);
  _G2.resize(n);
  auto rebuild = [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; i++) {
      _G1[i] = g1_from_body(b1.data() + i * G1_BYTES, 0);
      const bool checked = i < 2 || i + 1 == n || i == n / 2;
      _G2[i] = checked ? g2_from_body(b2.data() + i * G2_BYTES, 0) : g2_from_device_body(b2.data() + i * G2_BYTES);
    }
  };
  {
    unsigned hw = std::thread::hardware_concurrency();
    const size_t nt = n < 2048 ? 1 : std::min<size_t>(hw ? hw : 4, 16);
    std::vector<std::thread> pool;
    for (size_t t = 1; t < nt; t++) pool.emplace_back(rebuild, n * t / nt, n * (t + 1) / nt);
    rebuild(0, n / nt);
    for (auto& th : pool) th.join();
  }
  const Digest digest = digest_of(b1, b2), raw = raw_digest_of(_G1, _G2);
  std::lock_guard<std::mutex> lock(g_mu);
  remember_locked(this, side, digest, stamp_of(_G1, _G2));
  g_by_raw[raw] = ByRaw{side, digest};
}

// trusted_setup(filename): same file format and the same three failure modes as src/trusted_setup.cpp:76-121.
kzg::trusted_setup::trusted_setup(const std::string& filename) {
  kzg::init();
  std::ifstream file(filename, std::ios::in | std::ios::binary);
  if (!file.is_open()) throw runtime_error("could not open trusted setup file");
  std::vector<char> raw((std::istreambuf_iterator<char>(file)), std::istreambuf_iterator<char>());
  size_t pos = 0;
  auto take = [&](void* dst, size_t len) {
    if (pos + len > raw.size()) return false;
    memcpy(dst, raw.data() + pos, len);
    pos += len;
    return true;
  };
  uint64_t n = 0;
  if (!take(&n, sizeof n)) throw runtime_error("bad trusted setup file");
  _G1.reserve((size_t)n);
  _G2.reserve((size_t)n);
  for (int group = 1; group <= 2; group++) {
    const size_t cap = (group == 1 ? G1_BYTES : G2_BYTES) + 1;
    for (uint64_t i = 0; i < n; i++) {
      uint32_t len = 0;
      char body[G2_BYTES + 1];
      bool ok = take(&len, sizeof len) && len <= cap && take(body, len);
      octet o = {(int)len, (int)cap, body};
      if (group == 1) {
        ECP p;
        if (!ok || !ECP_fromOctet(&p, &o)) throw runtime_error("bad trusted setup file");
        _G1.push_back(p);
      } else {
        ECP2 p;
        if (!ok || !ECP2_fromOctet(&p, &o)) throw logic_error("bad trusted setup file");
        _G2.push_back(p);
      }
    }
  }
}

// Declared private in kzg.h; the reference's thread fan-out over index ranges is replaced by one device
// batch in the constructor above, so nothing calls this any more.
void kzg::trusted_setup::generate_elements_range(int, int, const std::vector<BIG>&) {
  throw logic_error("generate_elements_range is not used by the GPU trusted setup");
}

ECP kzg::trusted_setup::polyeval_G1(const ZZ_pX& P) {
  if (deg(P) == -1) return g1_from_body(nullptr, 1);
  SidePtr side = device_of(this, _G1, _G2);
  std::vector<uint8_t> raw = pack_coeffs(P);
  uint8_t xy[G1_BYTES];
  int is_inf = 0;
  // one device, or split by point range over all of them when the polynomial is large (kzg_gpu_group_msm_g1)
  int rc = kzg_gpu_group_msm_g1(side->group, raw.data(), (uint64_t)(deg(P) + 1), 0, xy, &is_inf);
  if (rc) throw std::runtime_error(std::string("kzg gpu: msm_g1: ") + kzg_gpu_group_last_error(side->group));
  return g1_from_body(xy, is_inf);
}

ECP2 kzg::trusted_setup::polyeval_G2(const ZZ_pX& P) {
  if (deg(P) == -1) return g2_from_body(nullptr, 1);
  SidePtr side = device_of(this, _G1, _G2);
  kzg_gpu_ctx* ctx = side->ctx();
  std::vector<uint8_t> raw = pack_coeffs(P);
  uint8_t xy[G2_BYTES];
  int is_inf = 0;
  int rc = kzg_gpu_msm_g2(ctx, raw.data(), (uint64_t)(deg(P) + 1), 0, xy, &is_inf);
  if (rc) fail(ctx, rc, "msm_g2");
  return g2_from_body(xy, is_inf);
}

kzg::commit kzg::trusted_setup::create_commit(const kzg::poly& poly) {
  if (deg(poly.get_poly()) + 1 >= (long)_G1.size())
    throw invalid_argument("polynomial degree be at most one less than the setup size (num_coeffs)");
  return kzg::commit(polyeval_G1(poly.get_poly()));
}

bool kzg::trusted_setup::verify_commit(kzg::commit& commit, const kzg::poly& poly) {
  kzg::commit again = create_commit(poly);
  return ECP_equals(&commit.get_curve_point(), &again.get_curve_point());
}

kzg::proof kzg::trusted_setup::create_proof(const kzg::poly& poly, int byte_offset, int byte_length, int chunk_size) {
  if (chunk_size > MAX_CHUNK_BYTES) throw invalid_argument("chunk_size must at most MAX_CHUNK_BYTES.");
  if (byte_offset % chunk_size != 0) throw invalid_argument("byte_offset is not a multiple of chunk_size.");
  if (byte_length % chunk_size != 0) throw invalid_argument("byte_length is not a multiple of chun_size.");
  return create_proof(poly, byte_offset / chunk_size, byte_length / chunk_size);
}

// q = P div prod (x - (chunk_offset + j)) and the MSM over q run back to back on the device
kzg::proof kzg::trusted_setup::create_proof(const kzg::poly& poly, int chunk_offset, int chunk_length) {
  if (chunk_length < 1) throw invalid_argument("chunk_length must be 1 or greater");
  const ZZ_pX& P = poly.get_poly();
  const long n = deg(P) + 1;
  if (n <= chunk_length) return kzg::proof(g1_from_body(nullptr, 1));  // zero quotient
  SidePtr side = device_of(this, _G1, _G2);
  kzg_gpu_ctx* ctx = side->ctx();
  std::vector<uint8_t> raw = pack_coeffs(P);
  uint8_t xy[G1_BYTES];
  int is_inf = 0;
  int rc = kzg_gpu_prove(ctx, raw.data(), (uint64_t)n, (int64_t)chunk_offset, (uint64_t)chunk_length, xy, &is_inf);
  if (rc) fail(ctx, rc, "prove");
  return kzg::proof(g1_from_body(xy, is_inf));
}

bool kzg::trusted_setup::verify_proof(kzg::commit& commit, kzg::proof& proof, kzg::blob& expected_data) {
  vector<pair<ZZ_p, ZZ_p>>& points = expected_data.get_data();
  if (points.size() < 1) throw invalid_argument("expected_data size must be 1 or greater");
  if (points.size() >= _G1.size()) return false;

  ZZ_pX interp, vanish;
  int64_t first_x = 0;
  std::vector<uint8_t> ys;
  if (kzg_b200::consecutive_nodes(points, first_x, ys))
    kzg_b200::interpolate_consecutive(first_x, ys, &interp, &vanish);  // device
  else
    linear_roots_and_polyfit(interp, vanish, points);                  // arbitrary nodes: the reference's host routine

  ECP2 z_s = polyeval_G2(vanish);                     // device
  ECP i_s = polyeval_G1(interp);                      // device
  // The reference compares fexp(ate(Z2, proof)) with fexp(ate(G2, commit - I1)) (src/trusted_setup.cpp:242-253): two
  // Miller loops and two final exponentiations.  Same predicate, one shared Miller loop and ONE final
  // exponentiation: e(Z2, proof) * e(G2, I1 - commit) == 1.
  ECP d = commit.get_curve_point();
  ECP_neg(&d);
  ECP_add(&d, &i_s);                                  // [I(s)]_1 - commit
  FP12 acc;
  PAIR_double_ate(&acc, &z_s, &proof.get_curve_point(), &_G2[0], &d);
  PAIR_fexp(&acc);
  return FP12_isunity(&acc);
}

void kzg::trusted_setup::export_setup(const std::string& filename) {
  std::ofstream file(filename, std::ios::out | std::ios::binary | std::ios::trunc);
  if (!file.is_open()) {
    std::cerr << "failed to export" << std::endl;
    return;
  }
  const uint64_t n = _G1.size();
  std::vector<char> out;
  out.reserve(8 + n * (size_t)(G1_BYTES + G2_BYTES + 10));
  auto put = [&](const void* p, size_t len) { out.insert(out.end(), (const char*)p, (const char*)p + len); };
  put(&n, sizeof n);
  for (uint64_t i = 0; i < n; i++) {
    char body[G1_BYTES + 1];
    octet o = {0, (int)sizeof(body), body};
    ECP_toOctet(&o, &_G1[i], false);
    const uint32_t len = (uint32_t)o.len;
    put(&len, sizeof len);
    put(o.val, o.len);
  }
  for (uint64_t i = 0; i < n; i++) {
    char body[G2_BYTES + 1];
    octet o = {0, (int)sizeof(body), body};
    ECP2_toOctet(&o, &_G2[i], false);
    const uint32_t len = (uint32_t)o.len;
    put(&len, sizeof len);
    put(o.val, o.len);
  }
  file.write(out.data(), (std::streamsize)out.size());
}
