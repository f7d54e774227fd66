"""workload for an ncu capture of direct_accumulate (ncu -k regex:direct_accumulate -c 1 --set full): 1024 blobs x 4096 coefficients
through the direct-lookup table (40 GiB grant).    python tools/ncu_direct.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
sys.path.insert(0, ROOT)
import kzg_b200  # noqa: E402
from bench import SECRET, random_scalars  # noqa: E402

ctx = kzg_b200.Context("bls12381", 0)
ctx.set_option("setup_g2", 0)
ctx.set_option("direct_table_gb", 40)
ctx.setup_powers(SECRET, 4097, want_bytes=False)
raw = random_scalars(1024 * 4096, 5).view(np.uint8).tobytes()
ctx.msm_g1_batch(raw, 4096, 1024)
ctx.close()
