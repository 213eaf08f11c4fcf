"""libyafaray_b200/data/noise_tables.bin: the lattice tables of the reference's noise bases (src/texture/noise_generator.cc:25-320 — hashpntf__, hashvectf__,
stdp_g__, hash__, stdp_p__; file-static literal data there) in the byte layout of yafgpu_scene_desc.noise_tables.  Dumped from the reference's own source by
oracle/ref_driver/noise_dump.cc (built by `make -C oracle _ref/gen/noise_tables.inc`, which also writes the header the C++ adaptor compiles in).
Needs /root/reference; the blob is committed like the other fixtures dumped from the reference (data/cie_xyz_360_830.npy)."""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "-B", "_ref/gen/noise_tables.inc"], check=True)
src = os.path.join(ROOT, "oracle", "_ref", "gen", "noise_tables.bin")
assert os.path.getsize(src) == 13340
shutil.copyfile(src, os.path.join(ROOT, "libyafaray_b200", "data", "noise_tables.bin"))
print("wrote libyafaray_b200/data/noise_tables.bin")
