// Thread-safety and side-table lifetime of the kzg.h drop-in (src/trusted_setup_gpu.cpp), built by
// src/build_dropin.sh against the reference's unmodified kzg.h and run by tests/test_gpu_dropin.py on the GPU box.
// The reference's create_commit / create_proof / verify_* only read _G1/_G2 and may be called concurrently once
// each thread has called kzg::init() (SURVEY.md 8(b) "Threading"); the drop-in must keep that property:
//   1. many threads making the FIRST use of one file-loaded setup at the same moment (one upload, no use-after-free);
//   2. more live setups than the resident-SRS cache holds, used round-robin from several threads (eviction while
//      another thread is inside a call on the evicted context);
//   3. copies of a setup share the resident SRS and give the same bytes.
#include <kzg.h>

#include <atomic>
#include <cstdio>
#include <string>
#include <thread>
#include <vector>

using namespace std;

static string random_text(unsigned seed, int len) {
  string s(len, 'a');
  for (int i = 0; i < len; i++) { seed = seed * 1664525u + 1013904223u; s[i] = (char)('a' + (seed >> 24) % 26); }
  return s;
}

int main(int argc, char** argv) {
  kzg::init();
  const string path = argc > 1 ? argv[1] : "/tmp/dropin_threads_srs.bin";
  atomic<int> bad(0);
  {
    kzg::trusted_setup made(300);
    made.export_setup(path);
  }
  // 1. concurrent first use of a file-loaded setup
  kzg::trusted_setup loaded(path);
  const int T = 8;
  vector<vector<uint8_t>> commits(T), proofs(T);
  {
    vector<thread> pool;
    for (int t = 0; t < T; t++)
      pool.emplace_back([&, t] {
        kzg::init();
        for (int rep = 0; rep < 6; rep++) {
          kzg::blob b = kzg::blob::from_string(random_text(1000 + t, 200));
          kzg::poly p = kzg::poly::from_blob(b);
          kzg::commit c = loaded.create_commit(p);
          kzg::proof pr = loaded.create_proof(p, 3 + t, 5);
          vector<uint8_t> cs = c.serialize(), ps = pr.serialize();
          if (rep == 0) { commits[t] = cs; proofs[t] = ps; }
          else if (cs != commits[t] || ps != proofs[t]) bad++;
          if (!loaded.verify_commit(c, p)) bad++;
          kzg::blob opened = kzg::blob::from_string(random_text(1000 + t, 200).substr(3 + t, 5), 3 + t);
          if (!loaded.verify_proof(c, pr, opened)) bad++;
        }
      });
    for (auto& th : pool) th.join();
  }
  // the same bytes single-threaded, through a COPY of the setup (shares the resident SRS)
  kzg::trusted_setup copy = loaded;
  for (int t = 0; t < T; t++) {
    kzg::poly p = kzg::poly::from_blob(kzg::blob::from_string(random_text(1000 + t, 200)));
    if (copy.create_commit(p).serialize() != commits[t]) bad++;
    if (copy.create_proof(p, 3 + t, 5).serialize() != proofs[t]) bad++;
  }
  // 2. more live setups than the cache holds (KZG_GPU_CACHE, default 8), used from several threads at once
  vector<kzg::trusted_setup> many;
  for (int i = 0; i < 11; i++) many.emplace_back(64 + i);
  {
    vector<thread> pool;
    for (int t = 0; t < 4; t++)
      pool.emplace_back([&, t] {
        kzg::init();
        for (int round = 0; round < 2; round++)
          for (int i = 0; i < (int)many.size(); i++) {
            const int k = (i + 3 * t) % (int)many.size();
            kzg::poly p = kzg::poly::from_blob(kzg::blob::from_string(random_text(77 + k, 40)));
            kzg::commit c = many[k].create_commit(p);
            if (!many[k].verify_commit(c, p)) bad++;
            kzg::proof pr = many[k].create_proof(p, 2, 3);
            kzg::blob opened = kzg::blob::from_string(random_text(77 + k, 40).substr(2, 3), 2);
            if (!many[k].verify_proof(c, pr, opened)) bad++;
            kzg::blob wrong = kzg::blob::from_string(random_text(78 + k, 40).substr(2, 3), 2);
            if (many[k].verify_proof(c, pr, wrong)) bad++;
          }
      });
    for (auto& th : pool) th.join();
  }
  printf("%s %d\n", bad.load() == 0 ? "OK" : "FAILED", bad.load());
  return bad.load() == 0 ? 0 : 1;
}
