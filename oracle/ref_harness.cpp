// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin command-line driver around the UNMODIFIED reference library
// (uncommitted6453/kzg-commitments, built from /root/reference by oracle/build_ref.sh).
// It exposes the reference's hot path (src/trusted_setup.cpp:137-254) through flat
// byte files so that tests/ and bench.py can (a) generate known-answer vectors and
// (b) time the reference's CPU path on the same inputs the CUDA path gets.
//
// Scalars/coefficients are flat little-endian records of SB = kzg::CURVE_ORDER_BYTES
// bytes each (exactly BytesFromZZ(buf, rep(c), SB); src/util.cpp:31-43 is the
// reference's own per-scalar marshalling).  Points are the reference's own
// serialisation (src/util.cpp:78-94): uint32 len | 0x04 | x | y, big-endian.
//
// `#define private public` is used only to reach polyeval_G1/polyeval_G2
// (src/kzg.h:187-188), which are the exact functions the CUDA MSM replaces.
#define private public
#include <kzg.h>
#undef private
#include "util.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

using namespace std;
using clk = std::chrono::steady_clock;

static vector<uint8_t> read_file(const string& path) {
  ifstream f(path, ios::binary);
  if (!f) { fprintf(stderr, "cannot open %s\n", path.c_str()); exit(2); }
  return vector<uint8_t>((istreambuf_iterator<char>(f)), istreambuf_iterator<char>());
}

static void write_file(const string& path, const vector<uint8_t>& b) {
  ofstream f(path, ios::binary | ios::trunc);
  f.write(reinterpret_cast<const char*>(b.data()), b.size());
}

// flat LE scalars -> ZZ_pX (not normalised away: NTL normalises, like the reference's polys)
static ZZ_pX poly_from_scalars(const vector<uint8_t>& raw) {
  const int SB = kzg::CURVE_ORDER_BYTES;
  long n = raw.size() / SB;
  ZZ_pX P;
  P.SetLength(n);
  for (long i = 0; i < n; i++) {
    ZZ z;
    ZZFromBytes(z, raw.data() + i * SB, SB);
    conv(P[i], z);
  }
  P.normalize();
  return P;
}

static vector<uint8_t> scalars_from_poly(const ZZ_pX& P, long n_out) {
  const int SB = kzg::CURVE_ORDER_BYTES;
  vector<uint8_t> out(n_out * SB, 0);
  for (long i = 0; i <= deg(P) && i < n_out; i++)
    BytesFromZZ(out.data() + i * SB, rep(P[i]), SB);
  return out;
}

static vector<uint8_t> serialize_ECP2(ECP2& Q) {
  // same framing as export_setup (src/trusted_setup.cpp:276-284)
  char buffer[4 * MODBYTES_CURVE + 1];
  octet oct = {0, (int)sizeof(buffer), buffer};
  ECP2_toOctet(&oct, &Q, false);
  vector<uint8_t> r;
  uint32_t len = oct.len;
  r.insert(r.end(), (uint8_t*)&len, (uint8_t*)&len + 4);
  r.insert(r.end(), (uint8_t*)oct.val, (uint8_t*)oct.val + oct.len);
  return r;
}

static double ms_since(clk::time_point t0) {
  return chrono::duration<double, milli>(clk::now() - t0).count();
}

static int usage() {
  fprintf(stderr,
    "usage: kzgref-<curve> <cmd> ...\n"
    "  info\n"
    "  setup <n> <srs_out>\n"
    "  commit <srs> <coeffs> <out>            create_commit -> commit.serialize()\n"
    "  prove <srs> <coeffs> <off> <k> <out>   create_proof(poly,off,k) -> proof.serialize()\n"
    "  verify <srs> <commit> <proof> <off> <ys>  verify_proof on blob (x=off+i, y=ys[i]) -> prints 0/1\n"
    "  g2msm <srs> <coeffs> <out>             polyeval_G2 -> framed ECP2 octet\n"
    "  quotient <coeffs> <off> <k> <out>      q=(P-I)/Z as the reference computes it\n"
    "  eval <coeffs> <off> <k> <out>          evaluate_polynomial_points\n"
    "  vanishing <off> <k> <out>              Z from linear_roots_and_polyfit (k+1 scalars)\n"
    "  interp <off> <ys> <out>                I from linear_roots_and_polyfit (k scalars)\n"
    "  polyser <coeffs> <out>                 poly.serialize()\n"
    "  time <srs> <coeffs> <what> <reps> [off k]   what=commit|prove|g2msm ; prints ms per rep\n");
  return 2;
}

int main(int argc, char** argv) {
  if (argc < 2) return usage();
  kzg::init();
  string cmd = argv[1];
  const int SB = kzg::CURVE_ORDER_BYTES;

  if (cmd == "info") {
    printf("{\"modbytes\": %d, \"scalar_bytes\": %d}\n", (int)MODBYTES_CURVE, SB);
    return 0;
  }
  if (cmd == "setup" && argc == 4) {
    kzg::trusted_setup ts(atoi(argv[2]));
    ts.export_setup(argv[3]);
    return 0;
  }
  if (cmd == "commit" && argc == 5) {
    kzg::trusted_setup ts{string(argv[2])};
    kzg::poly P(poly_from_scalars(read_file(argv[3])));
    kzg::commit c = ts.create_commit(P);
    write_file(argv[4], c.serialize());
    return 0;
  }
  if (cmd == "prove" && argc == 7) {
    kzg::trusted_setup ts{string(argv[2])};
    kzg::poly P(poly_from_scalars(read_file(argv[3])));
    kzg::proof pr = ts.create_proof(P, atoi(argv[4]), atoi(argv[5]));
    write_file(argv[6], pr.serialize());
    return 0;
  }
  if (cmd == "verify" && argc == 7) {
    kzg::trusted_setup ts{string(argv[2])};
    kzg::commit c = kzg::commit::deserialize(read_file(argv[3]));
    kzg::proof pr = kzg::proof::deserialize(read_file(argv[4]));
    int off = atoi(argv[5]);
    vector<uint8_t> ys = read_file(argv[6]);
    vector<pair<ZZ_p, ZZ_p>> pts;
    for (size_t i = 0; i < ys.size() / SB; i++) {
      ZZ z; ZZFromBytes(z, ys.data() + i * SB, SB);
      ZZ_p x, y; x = (long)(off + i); y = conv<ZZ_p>(z);
      pts.push_back({x, y});
    }
    kzg::blob b(pts);
    printf("%d\n", ts.verify_proof(c, pr, b) ? 1 : 0);
    return 0;
  }
  if (cmd == "g2msm" && argc == 5) {
    kzg::trusted_setup ts{string(argv[2])};
    ECP2 Q = ts.polyeval_G2(poly_from_scalars(read_file(argv[3])));
    write_file(argv[4], serialize_ECP2(Q));
    return 0;
  }
  if ((cmd == "quotient" || cmd == "eval") && argc == 6) {
    ZZ_pX P = poly_from_scalars(read_file(argv[2]));
    int off = atoi(argv[3]), k = atoi(argv[4]);
    vector<pair<ZZ_p, ZZ_p>> points;
    evaluate_polynomial_points(points, P, off, k);
    if (cmd == "eval") {
      vector<uint8_t> out(k * SB, 0);
      for (int i = 0; i < k; i++) BytesFromZZ(out.data() + i * SB, rep(points[i].second), SB);
      write_file(argv[5], out);
      return 0;
    }
    ZZ_pX I, Z;
    linear_roots_and_polyfit(I, Z, points);
    ZZ_pX q = (P - I) / Z;                      // src/trusted_setup.cpp:225
    long nq = deg(P) + 1 - k; if (nq < 0) nq = 0;
    write_file(argv[5], scalars_from_poly(q, nq));
    return 0;
  }
  if (cmd == "vanishing" && argc == 5) {
    int off = atoi(argv[2]), k = atoi(argv[3]);
    vector<pair<ZZ_p, ZZ_p>> pts;
    for (int i = 0; i < k; i++) { ZZ_p x, y; x = (long)(off + i); y = 0; pts.push_back({x, y}); }
    ZZ_pX I, Z;
    linear_roots_and_polyfit(I, Z, pts);
    write_file(argv[4], scalars_from_poly(Z, k + 1));
    return 0;
  }
  if (cmd == "interp" && argc == 5) {
    int off = atoi(argv[2]);
    vector<uint8_t> ys = read_file(argv[3]);
    vector<pair<ZZ_p, ZZ_p>> pts;
    for (size_t i = 0; i < ys.size() / SB; i++) {
      ZZ z; ZZFromBytes(z, ys.data() + i * SB, SB);
      ZZ_p x, y; x = (long)(off + i); y = conv<ZZ_p>(z);
      pts.push_back({x, y});
    }
    ZZ_pX I, Z;
    linear_roots_and_polyfit(I, Z, pts);
    write_file(argv[4], scalars_from_poly(I, pts.size()));
    return 0;
  }
  if (cmd == "polyser" && argc == 4) {
    kzg::poly P(poly_from_scalars(read_file(argv[2])));
    write_file(argv[3], P.serialize());
    return 0;
  }
  if (cmd == "time" && argc >= 6) {
    kzg::trusted_setup ts{string(argv[2])};
    ZZ_pX Pz = poly_from_scalars(read_file(argv[3]));
    kzg::poly P(Pz);
    string what = argv[4];
    int reps = atoi(argv[5]);
    int off = argc > 6 ? atoi(argv[6]) : 0, k = argc > 7 ? atoi(argv[7]) : 1;
    auto t0 = clk::now();
    for (int r = 0; r < reps; r++) {
      if (what == "commit") { kzg::commit c = ts.create_commit(P); (void)c; }
      else if (what == "prove") { kzg::proof p = ts.create_proof(P, off, k); (void)p; }
      else if (what == "g2msm") { ECP2 Q = ts.polyeval_G2(Pz); (void)Q; }
      else return usage();
    }
    printf("%.6f\n", ms_since(t0) / reps);
    return 0;
  }
  return usage();
}
