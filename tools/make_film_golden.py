"""Writes tests/golden/ref_cornell_32x18.film: a YAF_FILMv1 file saved by the reference's own ImageFilm::imageFilmSave
(oracle/_ref/yaf_ref_ieee with film_save_load = save), next to the raw accumulator dump of the same render (.npz).
The CPU tests pin libyafaray_b200/film_io.py against both."""
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import oracle_util as ou, parity_cases as pc  # noqa: E402
from libyafaray_b200 import film_io  # noqa: E402


def main():
    y = pc.film_io_scene(1, "save")
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "img.tga")
        raw, _ = ou.run_reference_render(y, film_path=out)
        shutil.copy(film_io.film_path(out, 3), os.path.join(ou.GOLDEN, "ref_cornell_32x18.film"))
        # the same folder resumed by the reference: one-pass film loaded, passes 2 and 3 of a three-pass schedule added (integrator_tiled.cc:198-202).
        # Not the uninterrupted three-pass render: a single-pass render places its samples differently (integrator_tiled.cc:392-403).
        resumed, _ = ou.run_reference_render(pc.film_io_scene(3, "load-save"), film_path=out)
    np.savez_compressed(os.path.join(ou.GOLDEN, "ref_cornell_32x18.npz"), film=raw)
    np.savez_compressed(os.path.join(ou.GOLDEN, "ref_cornell_32x18_resumed.npz"), film=resumed)
    print("written", raw.shape)


if __name__ == "__main__":
    main()
