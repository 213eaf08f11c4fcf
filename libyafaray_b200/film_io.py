"""The reference's on-disk film format ("YAF_FILMv1"): what ImageFilm::imageFilmSave writes and imageFilmLoad /
imageFilmLoadAllInFolder read (/root/reference/src/common/imagefilm.cc:1328-1465, 1467-1558, 1560-1657).  A film saved by a GPU render
can be merged or resumed by the reference's tools and the other way round (render farms mixing CPU and GPU nodes).

Layout (little endian, as File::append<T> writes raw host bytes, src/common/file.cc:190-201):
    "YAF_FILMv1\\0"                                   header string with its terminating NUL
    uint32 computer_node, base_sampling_offset, sampling_offset
    int32  w, h, cx0, cx1, cy0, cy1
    int32  n_image_passes, n_aux_image_passes
    per pass, per row y, per column x: float32 r, g, b, a, weight   — the [h, w, 5] accumulator layout of yafgpu_read_film

Only the combined pass (index 0) lives on the device; further passes of a loaded file are kept as arrays and written back unchanged."""
import os
import struct

import numpy as np

HEADER = b"YAF_FILMv1\x00"


class FilmFile:
    def __init__(self, passes, aux_passes=(), computer_node=0, base_sampling_offset=0, sampling_offset=0, cx0=0, cy0=0):
        self.passes = [np.ascontiguousarray(p, np.float32) for p in passes]
        self.aux_passes = [np.ascontiguousarray(p, np.float32) for p in aux_passes]
        self.computer_node, self.base_sampling_offset, self.sampling_offset = int(computer_node), int(base_sampling_offset), int(sampling_offset)
        self.cx0, self.cy0 = int(cx0), int(cy0)

    @property
    def film(self):
        return self.passes[0]

    @property
    def shape(self):
        return self.passes[0].shape[:2]


def film_path(image_output_path, computer_node=0):
    """ImageFilm::getFilmPath (imagefilm.cc:1328-1337): "<image output path> - node NNNN.film"."""
    return "%s - node %04d.film" % (image_output_path, computer_node)


def write_film_file(path, film, sampling_offset=0, base_sampling_offset=0, computer_node=0, cx0=0, cy0=0, extra_passes=(), aux_passes=()):
    """imageFilmSave.  `film` is the [h, w, 5] accumulator array of Context.read_film()."""
    film = np.ascontiguousarray(film, np.float32)
    if film.ndim != 3 or film.shape[2] != 5:
        raise ValueError("film must be [h, w, 5] float32 (r, g, b, a, weight)")
    h, w = film.shape[:2]
    passes = [film] + [np.ascontiguousarray(p, np.float32) for p in extra_passes]
    aux = [np.ascontiguousarray(p, np.float32) for p in aux_passes]
    for p in passes + aux:
        if p.shape != film.shape:
            raise ValueError("all passes must have the film's shape")
    with open(path, "wb") as fh:
        fh.write(HEADER)
        fh.write(struct.pack("<III", computer_node, base_sampling_offset, sampling_offset))
        fh.write(struct.pack("<iiiiii", w, h, cx0, cx0 + w, cy0, cy0 + h))   # cx_1_ = xstart + width, cy_1_ = ystart + height (imagefilm.cc:129-130)
        fh.write(struct.pack("<ii", len(passes), len(aux)))
        for p in passes + aux:
            fh.write(p.astype("<f4", copy=False).tobytes())
    return path


def read_film_file(path, expect_shape=None):
    """imageFilmLoad: returns a FilmFile; raises ValueError where the reference warns and refuses the file."""
    with open(path, "rb") as fh:
        data = fh.read()
    if not data.startswith(HEADER):
        raise ValueError("'%s' does not contain a valid YafaRay image film (header)" % path)
    o = len(HEADER)
    node, base, off = struct.unpack_from("<III", data, o); o += 12
    w, h, cx0, cx1, cy0, cy1 = struct.unpack_from("<iiiiii", data, o); o += 24
    n_img, n_aux = struct.unpack_from("<ii", data, o); o += 8
    if w <= 0 or h <= 0 or n_img < 1 or n_aux < 0 or cx1 - cx0 != w or cy1 - cy0 != h:
        raise ValueError("'%s': inconsistent film header" % path)
    if expect_shape is not None and tuple(expect_shape) != (h, w):
        raise ValueError("film check failed: image size expected %s, in loaded film %s" % (tuple(expect_shape), (h, w)))
    n = h * w * 5
    if len(data) - o != (n_img + n_aux) * n * 4:
        raise ValueError("'%s': truncated film (%d bytes of pixel data, expected %d)" % (path, len(data) - o, (n_img + n_aux) * n * 4))
    arr = np.frombuffer(data, "<f4", offset=o).reshape(n_img + n_aux, h, w, 5)
    return FilmFile([a.copy() for a in arr[:n_img]], [a.copy() for a in arr[n_img:]], node, base, off, cx0, cy0)


def load_films_in_folder(image_output_path, shape):
    """imageFilmLoadAllInFolder (imagefilm.cc:1467-1558): every "*.film" in the output folder whose file name starts with the output
    image's base name, in sorted order, accumulators summed, sampling offsets = the maximum.  Returns (FilmFile or None, loaded paths)."""
    d = os.path.dirname(image_output_path) or "."
    stem = os.path.splitext(os.path.basename(image_output_path))[0]
    paths = sorted(os.path.join(d, f) for f in os.listdir(d)
                   if os.path.isfile(os.path.join(d, f)) and f.endswith(".film") and f.startswith(stem))
    total, loaded = None, []
    for p in paths:
        try:
            f = read_film_file(p, shape)
        except ValueError:
            continue   # "Could not load film file": skipped with a warning
        if total is None:
            total = FilmFile([np.zeros_like(q) for q in f.passes], [np.zeros_like(q) for q in f.aux_passes])
        if len(f.passes) != len(total.passes) or len(f.aux_passes) != len(total.aux_passes):
            continue
        loaded.append(p)
        for acc, q in zip(total.passes + total.aux_passes, f.passes + f.aux_passes):
            acc += q   # float32 sums in file order, as the reference's loop
        total.sampling_offset = max(total.sampling_offset, f.sampling_offset)
        total.base_sampling_offset = max(total.base_sampling_offset, f.base_sampling_offset)
    return total, loaded
