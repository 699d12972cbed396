"""ctypes binding of libthorcuda.so (include/thorcuda.h).  Fails loudly when the library is absent."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

LF_DEBLOCK_Y, LF_DEBLOCK_UV, LF_CLPF, LF_MAKE_REF = 1, 2, 4, 8
LIST_PRED, LIST_TB, LIST_COST = 0, 1, 2
COST_SAD, COST_WIDESAD, COST_FASTHALF, COST_FASTQUARTER, COST_SSD = range(5)

# struct layouts of include/thorcuda.h
PRED_TASK = np.dtype([("x", "<u2"), ("y", "<u2"), ("w", "u1"), ("h", "u1"), ("ref0", "i1"), ("ref1", "i1"),
                      ("mv0x", "<i2"), ("mv0y", "<i2"), ("mv1x", "<i2"), ("mv1y", "<i2"),
                      ("sign0", "u1"), ("sign1", "u1"), ("dst", "u1"), ("reserved", "u1")])
TB_TASK = np.dtype([("x", "<u2"), ("y", "<u2"), ("size", "u1"), ("plane", "u1"), ("qp", "u1"), ("coeff_type", "u1"),
                    ("fast", "u1"), ("org", "u1"), ("pred", "u1"), ("rec", "u1"), ("coeff_offset", "<u4")])
COST_TASK = np.dtype([("x", "<u2"), ("y", "<u2"), ("w", "u1"), ("h", "u1"), ("org", "u1"), ("ref", "u1"), ("hx", "i1"), ("hy", "i1"),
                      ("ix", "<i2"), ("iy", "<i2"), ("reserved", "<u2")])
COST_RESULT = np.dtype([("cost", "<u4"), ("x", "<i2"), ("y", "<i2")])
SKIP_TASK = np.dtype([("x", "<u2"), ("y", "<u2"), ("size", "u1"), ("org", "u1"), ("pred", "u1"), ("reserved", "u1"),
                      ("thr_y", "<i4"), ("thr_c", "<i4")])
assert SKIP_TASK.itemsize == 16
INTRA_TASK = np.dtype([("x", "<u2"), ("y", "<u2"), ("cb_x", "<u2"), ("size", "u1"), ("plane", "u1"), ("mode", "u1"), ("flags", "u1"),
                       ("src", "u1"), ("dst", "u1")])
assert INTRA_TASK.itemsize == 12
MODE_DC, MODE_PLANAR, MODE_HOR, MODE_VER, MODE_UPLEFT, MODE_UPRIGHT, MODE_UPUPRIGHT, MODE_UPUPLEFT, MODE_UPLEFTLEFT, MODE_DOWNLEFTLEFT = range(10)
BLOCK_SYNTAX = np.dtype([(n, "u1") for n in ("frame_type", "num_ref", "enable_bipred", "interp_ref", "num_intra_modes", "max_delta_qp",
                                                "size", "ctx_index")] + [("ctx_cbp", "i1")] +
                        [(n, "u1") for n in ("max_num_pb_part", "max_num_tb_part", "num_skip_vec", "num_merge_vec")] + [("delta_qp", "i1")] +
                        [(n, "u1") for n in ("mode", "intra_mode", "pb_part", "tb_split", "skip_idx", "ref_idx0", "ref_idx1", "cbp_y", "cbp_u", "cbp_v")] +
                        [("mvp", "<i2", (2,)), ("mv0", "<i2", (4, 2)), ("mv1", "<i2", (4, 2)), ("tb_first", "<u4")])
assert BLOCK_SYNTAX.itemsize == 64
NO_TB = 0xffffffff
assert PRED_TASK.itemsize == 20 and TB_TASK.itemsize == 16 and COST_TASK.itemsize == 16 and COST_RESULT.itemsize == 8


class ThorCudaError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "libthorcuda.so")


def cdll():
    """The loaded library.  No fallback: a missing build is an error."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ThorCudaError(f"{p} not built: run `make` (or __graft_entry__.build()); there is no CPU fallback")
        _lib = C.CDLL(p)
        _lib.thorcuda_last_error.restype = C.c_char_p
        _lib.thorcuda_launch_count.restype = C.c_uint64
        _lib.thorcuda_stream.restype = C.c_void_p
        _lib.thorcuda_host_alloc.restype = C.c_void_p
        _lib.thorcuda_host_alloc.argtypes = [C.c_size_t]
        _lib.thorcuda_host_free.argtypes = [C.c_void_p]
        V, I, Z = C.c_void_p, C.c_int, C.c_size_t
        for name, res, args in (
                ("thorcuda_init", I, [I]),
                ("thorcuda_timer_begin", I, []), ("thorcuda_timer_end", I, [I]), ("thorcuda_timer_elapsed_ms", C.c_float, [I]),
                ("thorcuda_flush_l2", I, [Z]), ("thorcuda_calibrate_absdiff", C.c_double, [I]), ("thorcuda_fence_record", I, []), ("thorcuda_fence_wait", I, [I]),
                ("thorcuda_host_register", I, [V, Z]), ("thorcuda_host_unregister", None, [V]),
                ("thorcuda_pic_create", I, [I, I, I, I]), ("thorcuda_pic_destroy", None, [I]), ("thorcuda_pic_stride", I, [I, I]),
                ("thorcuda_pic_upload", I, [I, V, V, V, I, I]), ("thorcuda_pic_download", I, [I, V, V, V, I, I]),
                ("thorcuda_pic_download_padded", I, [I, V, V, V]), ("thorcuda_pic_upload_padded", I, [I, V, V, V]),
                ("thorcuda_pic_devptr", V, [I, I]), ("thorcuda_pic_copy", I, [I, I]),
                ("thorcuda_blockinfo_upload", I, [I, V]), ("thorcuda_blockinfo_upload_signals", I, [I, V, V]),
                ("thorcuda_loopfilter_frame", I, [I, I, I, I, I, V, V]), ("thorcuda_loopfilter_frames", I, [I, V]),
                ("thorcuda_pad_frame", I, [I]),
                ("thorcuda_list_create", I, [I, V, I, V, I]), ("thorcuda_list_destroy", None, [I]),
                ("thorcuda_list_pool_entries", Z, [I]),
                ("thorcuda_predict_list", I, [I, I, I]), ("thorcuda_predict_blocks", I, [V, I, V, I, I, I]),
                ("thorcuda_encode_list", I, [I, V, V, V]), ("thorcuda_encode_blocks", I, [V, I, V, I, V, V, V]),
                ("thorcuda_reconstruct_list", I, [I, V]), ("thorcuda_reconstruct_blocks", I, [V, I, V, I, V]),
                ("thorcuda_sad_table_build", I, [I, V, V, V, I, V]), ("thorcuda_sad_table_bytes", Z, [I, I]),
                ("thorcuda_sad_table_fetch", I, [I, I, V, Z]), ("thorcuda_sad_table_devptr", V, [I, I]),
                ("thorcuda_clip_mv", None, [V, I, I, I, I, I, I]),
                ("thorcuda_sad_table_index", C.c_int64, [I, I, I, I, V, I, I, I, I]),
                ("thorcuda_early_skip_thresholds", None, [I, I, C.c_float, V, V]),
                ("thorcuda_early_skip_batch", I, [V, I, V, I, V]),
                ("thorcuda_intra_predict_blocks", I, [V, I, V, I]),
                ("thorcuda_intra_reconstruct_levels", I, [V, I, V, V, V, V, I, V]),
                ("thorcuda_interpolate_frames", I, [I, I, I, I, I]), ("thorcuda_interp_levels", I, [I, I]),
                ("thorcuda_interp_field", I, [V, V, V]),
                ("thorcuda_coeff_bits_list", I, [I, V, V]), ("thorcuda_block_bits", I, [I, V, I, V]), ("thorcuda_block_bits_static", I, [V, I, V]), ("thorcuda_vlc_len", I, [C.c_uint, C.c_uint]),
                ("thorcuda_mv_bits", I, [I, I, I, I]),
                ("thorcuda_cost_list", I, [I, I, V]), ("thorcuda_cost_batch", I, [I, V, I, V, I, V])):
            fn = getattr(_lib, name)   # AttributeError here means a stale libthorcuda.so: rebuild
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc):
    if rc != 0:
        raise ThorCudaError(cdll().thorcuda_last_error().decode())


def init(device=0):
    check(cdll().thorcuda_init(int(device)))


def shutdown():
    if _lib is not None:
        _lib.thorcuda_shutdown()


def sync():
    check(cdll().thorcuda_sync())


def launch_count():
    return int(cdll().thorcuda_launch_count())


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Picture:
    """Device-resident I420 picture (handle into libthorcuda's pool); mirrors yuv_frame_t
    (reference common/types.h:41-59) + create_yuv_frame / close_yuv_frame."""

    def __init__(self, width, height, pad_y=0, pad_c=0):
        h = cdll().thorcuda_pic_create(width, height, pad_y, pad_c)
        if h < 0:
            raise ThorCudaError(cdll().thorcuda_last_error().decode())
        self.handle, self.width, self.height, self.pad_y, self.pad_c = h, width, height, pad_y, pad_c
        self.stride_y = cdll().thorcuda_pic_stride(h, 0)
        self.stride_c = cdll().thorcuda_pic_stride(h, 1)

    def close(self):
        if self.handle >= 0:
            cdll().thorcuda_pic_destroy(self.handle)
            self.handle = -1

    def upload(self, y, u, v):
        y, u, v = (np.ascontiguousarray(a, np.uint8) for a in (y, u, v))
        check(cdll().thorcuda_pic_upload(self.handle, _p(y), _p(u), _p(v), y.strides[0], u.strides[0]))
        sync()  # the numpy temporaries may die after return

    def download(self):
        y = np.empty((self.height, self.width), np.uint8)
        u = np.empty((self.height // 2, self.width // 2), np.uint8)
        v = np.empty_like(u)
        check(cdll().thorcuda_pic_download(self.handle, _p(y), _p(u), _p(v), self.width, self.width // 2))
        sync()
        return y, u, v

    def padded_shapes(self):
        return ((self.height + 2 * self.pad_y, self.stride_y), (self.height // 2 + 2 * self.pad_c, self.stride_c))

    def download_padded(self):
        sy, sc = self.padded_shapes()
        y, u, v = np.empty(sy, np.uint8), np.empty(sc, np.uint8), np.empty(sc, np.uint8)
        check(cdll().thorcuda_pic_download_padded(self.handle, _p(y), _p(u), _p(v)))
        sync()
        return y, u, v

    def upload_padded(self, y, u, v):
        sy, sc = self.padded_shapes()
        y, u, v = (np.ascontiguousarray(a, np.uint8) for a in (y, u, v))
        assert y.shape == sy and u.shape == sc and v.shape == sc
        check(cdll().thorcuda_pic_upload_padded(self.handle, _p(y), _p(u), _p(v)))
        sync()

    def devptr(self, plane):
        return cdll().thorcuda_pic_devptr(self.handle, plane)
