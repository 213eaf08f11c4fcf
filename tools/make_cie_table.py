"""libyafaray_b200/data/cie_xyz_360_830.npy: the CIE 1931 2-degree colour matching functions at 1 nm (471 rows of x, y, z) exactly as the
reference tabulates them, dumped from the reference itself (oracle/_ref/yaf_ref_ieee kat -> record "cie_xyz": wl2Xyz__ at the integer
wavelengths, oracle/ref_driver/kat.h).  Scientific data, not code; the device needs it for wl2RgbFromCie__ (glass dispersion)."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import oracle_util as ou  # noqa: E402

with tempfile.TemporaryDirectory() as td:
    p = os.path.join(td, "kat.bin")
    subprocess.check_call([ou.ref_bin(), "kat", p], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t = ou.read_kat(p)["cie_xyz"].reshape(471, 3).astype(np.float32)
out = os.path.join(ROOT, "libyafaray_b200", "data", "cie_xyz_360_830.npy")
os.makedirs(os.path.dirname(out), exist_ok=True)
np.save(out, t)
print(out, t.shape, "peak y at %d nm" % (360 + int(t[:, 1].argmax())), t[195])
