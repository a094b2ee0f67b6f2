"""ctypes binding of libmxrcnn_b200.so (include/mxrcnn_b200.h) + array plumbing.

There is NO CPU fallback: if the shared library is missing or no CUDA device is usable,
every operator raises.  Arrays cross the boundary as raw pointers:
  * numpy arrays      -> the ``*_host`` entry points (library stages H2D/D2H itself)
  * device arrays     -> anything exposing ``data_ptr()`` (torch), ``__cuda_array_interface__``
                         or ``__dlpack__`` (read with ctypes, no torch import needed)
"""
import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# MXRCNN_B200_LIB: another build of the same library (A/B measurements of kernel variants, tools/ab_build.py)
LIB_PATH = os.environ.get("MXRCNN_B200_LIB") or os.path.join(_HERE, "libmxrcnn_b200.so")

c_int_p = ctypes.POINTER(ctypes.c_int32)
c_float_p = ctypes.POINTER(ctypes.c_float)
c_double_p = ctypes.POINTER(ctypes.c_double)

MXR_DTYPE_F32, MXR_DTYPE_F64 = 0, 1


class MxrError(RuntimeError):
    pass


class ProposalParams(ctypes.Structure):
    _fields_ = [("batch", ctypes.c_int), ("num_anchors", ctypes.c_int), ("height", ctypes.c_int),
                ("width", ctypes.c_int), ("crop_h", ctypes.c_int), ("crop_w", ctypes.c_int),
                ("feat_stride", ctypes.c_double), ("pre_nms_top_n", ctypes.c_int), ("post_nms_top_n", ctypes.c_int),
                ("nms_thresh", ctypes.c_double), ("min_size", ctypes.c_float), ("first_stripe", ctypes.c_int)]


class AnchorTargetParams(ctypes.Structure):
    _fields_ = [("feat_height", ctypes.c_int), ("feat_width", ctypes.c_int), ("num_anchors", ctypes.c_int),
                ("feat_stride", ctypes.c_int), ("allowed_border", ctypes.c_int), ("im_height", ctypes.c_float),
                ("im_width", ctypes.c_float), ("num_gt", ctypes.c_int), ("negative_overlap", ctypes.c_double),
                ("positive_overlap", ctypes.c_double), ("clobber_positives", ctypes.c_int)]


class ProposalTargetParams(ctypes.Structure):
    _fields_ = [("num_rois", ctypes.c_int), ("num_gt", ctypes.c_int), ("num_keep", ctypes.c_int),
                ("fg_this", ctypes.c_int), ("num_classes", ctypes.c_int), ("normalize", ctypes.c_int),
                ("means", ctypes.c_double * 4), ("stds", ctypes.c_double * 4), ("inside_weights", ctypes.c_float * 4)]


class BoxHeadParams(ctypes.Structure):
    _fields_ = [("n_images", ctypes.c_int32), ("n_rois", ctypes.c_int32), ("n_classes", ctypes.c_int32),
                ("max_per_image", ctypes.c_int32), ("score_thresh", ctypes.c_double), ("nms_thresh", ctypes.c_double)]


# every symbol include/mxrcnn_b200.h declares: name -> (restype, argtypes)
_VP = ctypes.c_void_p
SIGNATURES = {
    "mxr_last_error_string": (ctypes.c_char_p, []),
    "mxr_version": (ctypes.c_int, []),
    "mxr_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(_VP)]),
    "mxr_destroy": (ctypes.c_int, [_VP]),
    "mxr_workspace_bytes": (ctypes.c_size_t, [_VP]),
    "mxr_synchronize": (ctypes.c_int, [_VP, _VP]),
    "mxr_profile_enable": (ctypes.c_int, [_VP, ctypes.c_int]),
    "mxr_profile_read": (ctypes.c_int, [_VP, ctypes.c_int, c_double_p, c_int_p]),
    "mxr_profile_stage_name": (ctypes.c_char_p, [ctypes.c_int]),
    "mxr_last_launch_count": (ctypes.c_int, [_VP]),
    "mxr_bench_alu_peak": (ctypes.c_int, [_VP, ctypes.c_int, c_double_p]),
    "mxr_proposal_phase_times": (ctypes.c_int, [_VP, ctypes.c_int, c_double_p, c_int_p]),
    "mxr_proposal_stamps": (ctypes.c_int, [_VP, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64), c_int_p]),
    "mxr_proposal_count_pairs": (ctypes.c_int, [_VP, ctypes.c_int]),
    "mxr_proposal_pairs_read": (ctypes.c_int, [_VP, ctypes.c_int, ctypes.POINTER(ctypes.c_uint32), c_int_p]),
    "mxr_decode_clip": (ctypes.c_int, [_VP, _VP, ctypes.POINTER(ProposalParams), _VP, _VP, c_double_p, _VP, _VP]),
    "mxr_topk_sorted": (ctypes.c_int, [_VP, _VP, ctypes.POINTER(ProposalParams), _VP, _VP, _VP, c_double_p, ctypes.c_int, _VP, _VP]),
    "mxr_proposal_counts": (ctypes.c_int, [_VP, ctypes.c_int, c_int_p, c_int_p]),
    "mxr_generate_anchors": (ctypes.c_int, [ctypes.c_double, c_double_p, ctypes.c_int, c_double_p, ctypes.c_int,
                                            c_double_p]),
    "mxr_reserve_proposal": (ctypes.c_int, [_VP, ctypes.POINTER(ProposalParams)]),
    "mxr_proposal_forward": (ctypes.c_int, [_VP, _VP, ctypes.POINTER(ProposalParams), _VP, _VP, _VP, c_double_p,
                                            ctypes.c_int, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_forward_host": (ctypes.c_int, [_VP, ctypes.POINTER(ProposalParams), _VP, _VP, _VP, c_double_p,
                                                 ctypes.c_int, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_submit_host": (ctypes.c_int, [_VP, ctypes.c_int, ctypes.POINTER(ProposalParams), _VP, _VP, _VP,
                                                c_double_p, ctypes.c_int]),
    "mxr_proposal_submit_host_packed": (ctypes.c_int, [_VP, ctypes.c_int, ctypes.POINTER(ProposalParams), _VP, c_double_p,
                                                       ctypes.c_int]),
    "mxr_proposal_wait_host": (ctypes.c_int, [_VP, ctypes.c_int, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_pad": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_int_p, ctypes.c_int,
                                        _VP, _VP]),
    "mxr_nms": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, _VP, _VP]),
    "mxr_nms_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, _VP, _VP]),
    "mxr_nms_batched": (ctypes.c_int, [_VP, _VP, _VP, c_int_p, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                       ctypes.c_int, _VP, _VP]),
    "mxr_nms_batched_host": (ctypes.c_int, [_VP, _VP, c_int_p, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int,
                                            _VP, _VP]),
    "mxr_nest": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, _VP]),
    "mxr_nest_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, _VP]),
    "mxr_box_head": (ctypes.c_int, [_VP, _VP, ctypes.POINTER(BoxHeadParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_box_head_host": (ctypes.c_int, [_VP, ctypes.POINTER(BoxHeadParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_bbox_pred": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, _VP]),
    "mxr_bbox_pred_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, _VP]),
    "mxr_clip_boxes": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                      ctypes.c_double]),
    "mxr_clip_boxes_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                           ctypes.c_double]),
    "mxr_bbox_transform": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int, _VP]),
    "mxr_bbox_transform_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int, _VP]),
    "mxr_bbox_overlaps": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int,
                                         _VP, _VP, _VP, _VP]),
    "mxr_bbox_overlaps_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int,
                                              _VP, _VP, _VP, _VP]),
    "mxr_expand_bbox_targets": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, c_float_p, _VP, _VP]),
    "mxr_expand_bbox_targets_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, c_float_p, _VP, _VP]),
    "mxr_anchor_target_phase1": (ctypes.c_int, [_VP, _VP, c_double_p, _VP, c_int_p]),
    "mxr_anchor_target_phase2": (ctypes.c_int, [_VP, _VP, c_int_p, ctypes.c_int, c_int_p, ctypes.c_int, ctypes.c_double,
                                                ctypes.c_double, c_float_p, _VP, _VP, _VP, _VP]),
    "mxr_anchor_target_batch_phase1": (ctypes.c_int, [_VP, _VP, c_double_p, ctypes.c_int, c_float_p, _VP, c_int_p, c_int_p]),
    "mxr_anchor_target_batch_phase2": (ctypes.c_int, [_VP, _VP, c_int_p, c_int_p, c_double_p, c_float_p, _VP, _VP, _VP, _VP]),
    "mxr_anchor_target_phase2_host": (ctypes.c_int, [_VP, c_int_p, ctypes.c_int, c_int_p, ctypes.c_int, ctypes.c_double,
                                                     ctypes.c_double, c_float_p, _VP, _VP, _VP, _VP]),
    "mxr_recall_cover_host": (ctypes.c_int, [_VP, c_double_p, ctypes.c_int, c_double_p, ctypes.c_int, c_double_p]),
    "mxr_fg_bg_lists": (ctypes.c_int, [_VP, _VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                       ctypes.c_double, _VP, _VP, _VP]),
    "mxr_fg_bg_lists_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                            ctypes.c_double, _VP, _VP, _VP]),
    "mxr_proposal_target_gather": (ctypes.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_target_gather_host": (ctypes.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_target_phase1_host": (ctypes.c_int, [_VP, _VP, ctypes.c_int, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                       ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double,
                                                       ctypes.c_double, _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_proposal_target_phase2_host": (ctypes.c_int, [_VP, ctypes.POINTER(ProposalTargetParams), _VP, _VP, _VP, _VP, _VP, _VP]),
    "mxr_nms_noalloc": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, _VP, _VP]),
}

_lib = None
_lib_lock = threading.Lock()


def load():
    """Load the shared library (building it is __graft_entry__.build()'s / mxrcnn_b200.build's job)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise MxrError("libmxrcnn_b200.so is not built (run `python -m mxrcnn_b200.build`); "
                               "there is no CPU fallback")
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                if os.environ.get("MXRCNN_B200_LIB") and not hasattr(lib, name):
                    continue  # an older A/B build may lack the newest entry points
                fn = getattr(lib, name)
                fn.restype, fn.argtypes = res, args
            _lib = lib
    return _lib


_STATUS_EXC = {-1: ValueError, -2: MxrError, -3: ValueError, -4: ValueError, -5: MemoryError, -6: ValueError}


def check(status):
    if status != 0:
        msg = load().mxr_last_error_string().decode("utf-8", "replace")
        raise _STATUS_EXC.get(status, MxrError)(msg)


class Handle(object):
    """One per (device, thread).  Owns the library workspace."""

    def __init__(self, device=0):
        lib = load()
        self._h = _VP()
        check(lib.mxr_create(int(device), ctypes.byref(self._h)))
        self.device = int(device)

    @property
    def ptr(self):
        return self._h

    def synchronize(self, stream=None):
        check(load().mxr_synchronize(self._h, _VP(stream or 0)))

    def profile_enable(self, on=True):
        check(load().mxr_profile_enable(self._h, 1 if on else 0))

    def profile_read(self):
        """{kernel name: (total ms, launches)} since the last read; synchronises the device."""
        n = 32
        ms = (ctypes.c_double * n)()
        cnt = (ctypes.c_int32 * n)()
        check(load().mxr_profile_read(self._h, n, ms, cnt))
        lib = load()
        return {lib.mxr_profile_stage_name(i).decode(): (ms[i], cnt[i]) for i in range(n) if cnt[i]}

    PHASES = ("startup_barrier", "decode_filter", "radix_select", "compact", "head_sort", "redecode", "nms_stripes",
              "outputs")

    def proposal_phase_times(self, max_images=64):
        """Per-image phase durations (us) of the last fused proposal launch: list of dicts."""
        us = (ctypes.c_double * (max_images * 9))()
        n = ctypes.c_int32(0)
        check(load().mxr_proposal_phase_times(self._h, max_images, us, ctypes.byref(n)))
        out = []
        for b in range(n.value):
            d = {name: us[b * 9 + i] for i, name in enumerate(self.PHASES)}
            d["stripes_walked"] = int(us[b * 9 + 8])
            out.append(d)
        return out

    def proposal_stamps(self, max_images=64):
        """Raw %globaltimer stamps (ns), counters and cycle stamps of the last fused proposal launch: (images, 48) uint64."""
        buf = np.zeros((max_images, 48), np.uint64)
        n = ctypes.c_int32(0)
        check(load().mxr_proposal_stamps(self._h, max_images, buf.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                         ctypes.byref(n)))
        return buf[:n.value]

    def proposal_counts(self, max_images=1024):
        """(images, 4) int32 of the last proposal call: n_valid, boxes entering NMS, boxes kept, NaN flag."""
        buf = np.zeros((max_images, 4), np.int32)
        n = ctypes.c_int32(0)
        check(load().mxr_proposal_counts(self._h, max_images, buf.ctypes.data_as(c_int_p), ctypes.byref(n)))
        return buf[:n.value]

    def count_pairs(self, on):
        check(load().mxr_proposal_count_pairs(self._h, 1 if on else 0))

    def pairs_read(self, max_images=1024):
        """(images, 2): IoU pairs evaluated / settled by the float64 test, per image, by the last fused proposal call
        made while count_pairs(True) was set."""
        buf = np.zeros((max_images, 2), np.uint32)
        n = ctypes.c_int32(0)
        check(load().mxr_proposal_pairs_read(self._h, max_images, buf.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)),
                                             ctypes.byref(n)))
        return buf[:n.value]

    def last_launch_count(self):
        return int(load().mxr_last_launch_count(self._h))

    def workspace_bytes(self):
        return int(load().mxr_workspace_bytes(self._h))

    def close(self):
        if self._h:
            load().mxr_destroy(self._h)
            self._h = _VP()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_tls = threading.local()


def default_handle(device=None):
    """Per-thread, per-device handle used by the module-level helper functions."""
    if device is None:
        device = int(os.environ.get("MXRCNN_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    cache = getattr(_tls, "handles", None)
    if cache is None:
        cache = _tls.handles = {}
    if device not in cache:
        cache[device] = Handle(device)
    return cache[device]


def device_id_of(x):
    """CUDA device ordinal a device array lives on."""
    if hasattr(x, "is_cuda"):
        idx = x.device.index
        if idx is None:
            import torch
            idx = torch.cuda.current_device()
        return int(idx)
    if hasattr(x, "__dlpack_device__"):
        return int(x.__dlpack_device__()[1])
    return 0


def handle_for(handle, *arrays):
    """The handle a helper should use: the caller's (checked against the device the arrays live on), else the
    calling thread's default handle of THAT device -- never a handle of another GPU than the data's."""
    dev = None
    for a in arrays:
        if a is not None and is_device_array(a):
            dev = device_id_of(a)
            break
    if handle is not None:
        if dev is not None and handle.device != dev:
            raise ValueError("handle of device %d used with arrays on device %d" % (handle.device, dev))
        return handle
    return default_handle(dev)


# ---------------------------------------------------------------------------------------
# array plumbing
# ---------------------------------------------------------------------------------------
class _DLDevice(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32)]


class _DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class _DLTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("device", _DLDevice), ("ndim", ctypes.c_int32), ("dtype", _DLDataType),
                ("shape", ctypes.POINTER(ctypes.c_int64)), ("strides", ctypes.POINTER(ctypes.c_int64)),
                ("byte_offset", ctypes.c_uint64)]


_DL_CODES = {(2, 32): np.float32, (2, 64): np.float64, (0, 32): np.int32, (0, 64): np.int64, (1, 8): np.uint8}


class DeviceView(object):
    """Raw view of a device array: pointer, shape, dtype.  Keeps its owner alive."""

    def __init__(self, ptr, shape, dtype, device_id, owner):
        self.ptr, self.shape, self.dtype, self.device_id, self._owner = int(ptr), tuple(shape), np.dtype(dtype), device_id, owner

    @property
    def size(self):
        return int(np.prod(self.shape)) if self.shape else 1


def is_device_array(x):
    if isinstance(x, np.ndarray) or x is None:
        return False
    if hasattr(x, "is_cuda"):
        return bool(x.is_cuda)
    if hasattr(x, "__cuda_array_interface__"):
        return True
    if hasattr(x, "__dlpack_device__"):
        return int(x.__dlpack_device__()[0]) == 2  # kDLCUDA
    return False


def device_view(x, dtype=None):
    """DeviceView of a C-contiguous device array (torch tensor, CAI object or DLPack producer)."""
    if hasattr(x, "data_ptr") and hasattr(x, "is_contiguous"):  # torch, without importing it
        if not x.is_contiguous():
            raise ValueError("device arrays must be C-contiguous")
        np_dt = np.dtype(str(x.dtype).replace("torch.", ""))
        v = DeviceView(x.data_ptr(), tuple(x.shape), np_dt, x.device.index or 0, x)
    elif hasattr(x, "__cuda_array_interface__"):
        cai = x.__cuda_array_interface__
        if cai.get("strides") is not None:
            raise ValueError("device arrays must be C-contiguous")
        v = DeviceView(cai["data"][0], cai["shape"], np.dtype(cai["typestr"]), 0, x)
    elif hasattr(x, "__dlpack__"):
        cap = x.__dlpack__()
        ctypes.pythonapi.PyCapsule_GetPointer.restype = ctypes.c_void_p
        ctypes.pythonapi.PyCapsule_GetPointer.argtypes = [ctypes.py_object, ctypes.c_char_p]
        t = ctypes.cast(ctypes.pythonapi.PyCapsule_GetPointer(cap, b"dltensor"), ctypes.POINTER(_DLTensor)).contents
        if t.device.device_type != 2:
            raise ValueError("DLPack tensor is not on a CUDA device")
        shape = [t.shape[i] for i in range(t.ndim)]
        if t.strides:
            exp = 1
            for i in reversed(range(t.ndim)):
                if shape[i] != 1 and t.strides[i] != exp:
                    raise ValueError("device arrays must be C-contiguous")
                exp *= shape[i]
        np_dt = _DL_CODES[(t.dtype.code, t.dtype.bits)]
        v = DeviceView((t.data or 0) + t.byte_offset, shape, np_dt, t.device.device_id, (x, cap))
    else:
        raise TypeError("not a device array: %r" % type(x))
    if dtype is not None and v.dtype != np.dtype(dtype):
        raise TypeError("expected a %s device array, got %s" % (np.dtype(dtype), v.dtype))
    return v


def current_stream_ptr(x=None):
    """CUDA stream the producer of ``x`` is working on (torch's current stream) or 0."""
    if x is not None and hasattr(x, "is_cuda"):
        import torch  # the caller already uses torch; the library itself does not need it
        return int(torch.cuda.current_stream(x.device).cuda_stream)
    return 0


def host(a, dtype):
    """C-contiguous numpy array of ``dtype`` (no copy when already so)."""
    return np.ascontiguousarray(a, dtype=dtype)


def hptr(a):
    return _VP(a.ctypes.data) if a is not None else _VP(0)
