"""ctypes driver of a libkvazaar build through its public API -- kvz_api_get() and the twelve entry points of
`struct kvz_api` (src/kvazaar.h:545-707) -- the way the reference's own CLI drives it (src/encmain.c:360-640).

It is used with kvazaar_b200/host/_build/libkvazaar_cuda.so (the reference encoder with the CUDA path bound in by
kvazaar_b200/host/, KVZ_CUDA_MODE=batch) by bench.py and the tests; the same class drives an unmodified libkvazaar for
the CPU baseline.  Nothing here computes: pictures go in, bitstream chunks come out.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_LIB = os.path.join(ROOT, "kvazaar_b200", "host", "_build", "libkvazaar_cuda.so")

KVZ_DATA_CHUNK_SIZE = 4096


class KvzPicture(C.Structure):
    """struct kvz_picture (src/kvazaar.h:394-417)"""


KvzPicture._fields_ = [
    ("fulldata", C.c_void_p), ("y", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("data", C.c_void_p * 3),
    ("width", C.c_int32), ("height", C.c_int32), ("stride", C.c_int32),
    ("base_image", C.POINTER(KvzPicture)), ("refcount", C.c_int32),
    ("pts", C.c_int64), ("dts", C.c_int64), ("interlacing", C.c_int), ("chroma_format", C.c_int),
    ("ref_pocs", C.c_int32 * 16),
]


class KvzDataChunk(C.Structure):
    """struct kvz_data_chunk (src/kvazaar.h:529-538)"""


KvzDataChunk._fields_ = [("data", C.c_uint8 * KVZ_DATA_CHUNK_SIZE), ("len", C.c_uint32), ("next", C.POINTER(KvzDataChunk))]


class KvzFrameInfo(C.Structure):
    """struct kvz_frame_info (src/kvazaar.h:492-527)"""
    _fields_ = [("poc", C.c_int32), ("qp", C.c_int8), ("nal_unit_type", C.c_int), ("slice_type", C.c_int),
                ("ref_list", (C.c_int * 16) * 2), ("ref_list_len", C.c_int * 2)]


_PIC_P = C.POINTER(KvzPicture)
_CHUNK_P = C.POINTER(KvzDataChunk)


class KvzApi(C.Structure):
    """struct kvz_api (src/kvazaar.h:545-704)"""
    _fields_ = [
        ("config_alloc", C.CFUNCTYPE(C.c_void_p)),
        ("config_destroy", C.CFUNCTYPE(C.c_int, C.c_void_p)),
        ("config_init", C.CFUNCTYPE(C.c_int, C.c_void_p)),
        ("config_parse", C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_char_p)),
        ("picture_alloc", C.CFUNCTYPE(_PIC_P, C.c_int32, C.c_int32)),
        ("picture_free", C.CFUNCTYPE(None, _PIC_P)),
        ("chunk_free", C.CFUNCTYPE(None, _CHUNK_P)),
        ("encoder_open", C.CFUNCTYPE(C.c_void_p, C.c_void_p)),
        ("encoder_close", C.CFUNCTYPE(None, C.c_void_p)),
        ("encoder_headers", C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(_CHUNK_P), C.POINTER(C.c_uint32))),
        ("encoder_encode", C.CFUNCTYPE(C.c_int, C.c_void_p, _PIC_P, C.POINTER(_CHUNK_P), C.POINTER(C.c_uint32),
                                       C.POINTER(_PIC_P), C.POINTER(_PIC_P), C.POINTER(KvzFrameInfo))),
        ("picture_alloc_csp", C.CFUNCTYPE(_PIC_P, C.c_int, C.c_int32, C.c_int32)),
    ]


class EncoderError(RuntimeError):
    pass


_SHORT = {"r": "ref", "q": "qp", "p": "period", "w": "width", "h": "height"}     # src/cli.c: short option -> long name


def options_from_cli(text):
    """'--preset medium --me full32 --gop 8 -r 4 --owf 0' -> ordered {long option: value or None}; -n / -i / -o and
    --input-res belong to the CLI, not to kvz_config, and are dropped"""
    out, toks, i = {}, text.split(), 0
    while i < len(toks):
        k = toks[i].lstrip("-")
        k = _SHORT.get(k, k)
        val = None
        if i + 1 < len(toks) and not (toks[i + 1].startswith("-") and not toks[i + 1].lstrip("-").replace(".", "").isdigit()):
            val = toks[i + 1]
            i += 1
        i += 1
        if k not in ("n", "frames", "i", "input", "o", "output", "input-res"):
            out[k] = val
    return out


class Encoder:
    """One open encoder ("Only one encoder may be open at a time", src/kvazaar.h:619).

    options: the CLI's long options without the leading dashes, e.g.
        {"preset": "medium", "me": "full32", "gop": "8", "ref": "4", "owf": "0", "threads": "16"}
    (the CLI applies a preset first, then the other options in command-line order, src/cli.c; so does this).
    """

    def __init__(self, lib_path, width, height, options, quiet=True):
        if not os.path.exists(lib_path):
            raise EncoderError(f"{lib_path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        # RTLD_GLOBAL: libkvz_cuda.so finds the host's kvz_strategyselector_register with dlsym(RTLD_DEFAULT, ...)
        self.lib = C.CDLL(lib_path, mode=C.RTLD_GLOBAL)
        self.lib.kvz_api_get.restype = C.POINTER(KvzApi)
        self.api = self.lib.kvz_api_get(8).contents
        self.width, self.height = width, height
        cfg = self.api.config_alloc()
        if not cfg or not self.api.config_init(cfg):
            raise EncoderError("kvz config_alloc / config_init failed")
        opts = dict(options)
        ordered = ([("preset", opts.pop("preset"))] if "preset" in opts else []) + list(opts.items())
        ordered = [("input-res", f"{width}x{height}")] + ordered
        for k, v in ordered:
            if not self.api.config_parse(cfg, k.encode(), b"true" if v is None else str(v).encode()):
                raise EncoderError(f"kvz config_parse rejected --{k} {v}")
        saved = None
        if quiet:                                   # the strategy selector reports to stderr
            devnull = os.open(os.devnull, os.O_WRONLY)
            saved = os.dup(2)
            os.dup2(devnull, 2)
            os.close(devnull)
        try:
            self.enc = self.api.encoder_open(cfg)
        finally:
            if saved is not None:
                os.dup2(saved, 2)
                os.close(saved)
        self.api.config_destroy(cfg)
        if not self.enc:
            raise EncoderError("kvz encoder_open failed (with the CUDA path bound in: no usable GPU, there is no CPU fallback)")
        self.frames_in = 0

    def close(self):
        if getattr(self, "enc", None):
            self.api.encoder_close(self.enc)
            self.enc = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def picture(self, y, u, v):
        """kvz_picture filled from three numpy planes (uint8, 4:2:0); freed by encode()."""
        pic = self.api.picture_alloc(self.width, self.height)
        if not pic:
            raise EncoderError("kvz picture_alloc failed")
        p = pic.contents
        for plane, ptr, (w, h) in ((y, p.y, (self.width, self.height)), (u, p.u, (self.width // 2, self.height // 2)),
                                   (v, p.v, (self.width // 2, self.height // 2))):
            stride = p.stride if plane is y else p.stride // 2
            dst = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(h, stride))
            dst[:, :w] = plane
        p.pts = self.frames_in
        self.frames_in += 1
        return pic

    def encode(self, pic):
        """One encoder_encode call (pic = None flushes).  Returns (bytes or None, frame_info or None): with --gop 8 the
        first calls buffer pictures and return nothing; with --owf 0 a call that returns data has encoded that picture."""
        chunks, length = _CHUNK_P(), C.c_uint32(0)
        rec, src = _PIC_P(), _PIC_P()
        info = KvzFrameInfo()
        ok = self.api.encoder_encode(self.enc, pic if pic is not None else _PIC_P(), C.byref(chunks), C.byref(length),
                                     C.byref(rec), C.byref(src), C.byref(info))
        if pic is not None:
            self.api.picture_free(pic)
        if not ok:
            raise EncoderError("kvz encoder_encode failed")
        data = None
        if chunks:
            parts, c = [], chunks
            while c:
                parts.append(bytes(c.contents.data[:c.contents.len]))
                c = c.contents.next
            data = b"".join(parts)
            self.api.chunk_free(chunks)
        if rec:
            self.api.picture_free(rec)
        if src:
            self.api.picture_free(src)
        return data, (info if data is not None else None)
