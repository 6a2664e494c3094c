"""Tensor-level entry points: checks + output allocation + current stream + C-ABI call.

These mirror the ten pybind functions of the reference's `pointnet2._ext`
(pointnet2/_ext_src/src/bindings.cpp:13-26) and `knn_pytorch.knn`
(knn/src/vision.cpp:3-5) -- same names, argument order and error behaviour
(non-contiguous / wrong dtype / non-CUDA input -> RuntimeError, as the
reference's CHECK_* macros in include/utils.h:10-30 do) -- plus the two
post-processing ops.  PyTorch is used for device memory and streams only.
"""
import ctypes
import math

import torch

from . import _abi

_lib = _abi.lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _chk(t, name, dtype):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError("%s must be a tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (b200grasp has no CPU path)" % name)
    if not t.is_contiguous():
        raise RuntimeError("%s must be a contiguous tensor" % name)
    if t.dtype != dtype:
        raise RuntimeError("%s must be a %s tensor" % (name, str(dtype).replace("torch.", "")))


def furthest_point_sampling(points, nsamples):
    """points (B,N,3) f32 -> (B,nsamples) i32   [sampling.cpp:70-91]"""
    _chk(points, "points", torch.float32)
    b, n, _ = points.shape
    out = torch.empty((b, nsamples), dtype=torch.int32, device=points.device)
    with torch.cuda.device(points.device):
        _abi.check(_lib.b200_furthest_point_sampling(_ptr(points), b, n, nsamples, _ptr(out), _stream()),
                   "furthest_point_sampling")
    return out


def fps_prefix_check(xyz, n, m):
    """Is FPS over the first n points of each cloud of xyz (B,Ntot,3), m samples, the identity permutation?
    -> verdict tensor (B,m) i32 whose column 0 is 0 (yes) or -1 (run the chain); pass it to
    furthest_point_sampling(..., verdicts=[...]).  No synchronisation."""
    _chk(xyz, "xyz", torch.float32)
    b, ntot, _ = xyz.shape
    if n > ntot:
        raise ValueError("b200grasp: prefix longer than the cloud")
    verdict = torch.empty((b, m), dtype=torch.int32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_fps_prefix_check(_ptr(xyz), b, n, m, ntot, _ptr(verdict), _stream()), "fps_prefix_check")
    return verdict


def furthest_point_sampling_spec(xyz, npoint, verdicts):
    """furthest_point_sampling whose identity shortcut is decided by verdict tensors from fps_prefix_check (this
    level's and its ancestors'): same result as furthest_point_sampling, always."""
    _chk(xyz, "xyz", torch.float32)
    if not 1 <= len(verdicts) <= 3:
        raise ValueError("b200grasp: one to three verdict tensors")
    b, n, _ = xyz.shape
    out = torch.empty((b, npoint), dtype=torch.int32, device=xyz.device)
    args = []
    for k in range(3):
        if k < len(verdicts):
            _chk(verdicts[k], "verdict", torch.int32)
            args += [_ptr(verdicts[k]), verdicts[k].shape[1]]
        else:
            args += [None, 0]
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_furthest_point_sampling_spec(_ptr(xyz), b, n, npoint, *args, _ptr(out), _stream()),
                   "furthest_point_sampling_spec")
    return out


def gather_points(points, idx):
    """points (B,C,N) f32, idx (B,M) i32 -> (B,C,M)   [sampling.cpp:20-43]"""
    _chk(points, "points", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, c, n = points.shape
    m = idx.shape[1]
    out = torch.empty((b, c, m), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        _abi.check(_lib.b200_gather_points(_ptr(points), _ptr(idx), b, c, n, m, _ptr(out), _stream()),
                   "gather_points")
    return out


def gather_xyz(xyz, idx, transposed=True):
    """xyz (B,N,3) f32, idx (B,M) i32 -> new_xyz (B,M,3) [, new_xyz_t (B,3,M)] in one launch
    (= gather_operation(xyz^T, idx) and its transpose, pointnet2_modules.py:230-240)"""
    _chk(xyz, "xyz", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, n, _ = xyz.shape
    m = idx.shape[1]
    out = torch.empty((b, m, 3), dtype=torch.float32, device=xyz.device)
    out_t = torch.empty((b, 3, m), dtype=torch.float32, device=xyz.device) if transposed else None
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_gather_xyz(_ptr(xyz), _ptr(idx), b, n, m, _ptr(out),
                                        _ptr(out_t) if transposed else None, _stream()), "gather_xyz")
    return (out, out_t) if transposed else out


def mlp_pack_weights(w):
    """w (Cout,Cin) f32 (1x1 conv weight with BatchNorm folded in) -> the hi | lo K-slab packing
    group_mlp_maxpool streams (2 * Cout * Cin floats)"""
    _chk(w, "w", torch.float32)
    cout, cin = w.shape
    packed = torch.empty(2 * cout * cin, dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        _abi.check(_lib.b200_mlp_pack_weights(_ptr(w), cout, cin, _ptr(packed), _stream()), "mlp_pack_weights")
    return packed


def group_mlp_maxpool(xyz, new_xyz, idx, layers, radius=0.0, rot=None, passes=3, channels_last=False):
    """Fused group -> (centre, /radius, rotate) -> 3 x (1x1 conv + folded BN + ReLU) -> max over nsample.
    xyz (B,N,3), new_xyz (B,M,3), idx (B,M,ns) or (G,B,M,ns) i32, rot (B,M,3,3)/(B,M,9) or None,
    layers = [(w1 (C1,3), b1), (w2_packed, b2, C2), (w3_packed, b3, C3)] -> (B,C3,M) or (B,C3,M,G)
    [, the same channels-last (B,M,C3) / (B,M,G,C3) when channels_last]."""
    _chk(xyz, "xyz", torch.float32)
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(idx, "idx", torch.int32)
    squeeze = idx.dim() == 3
    g = 1 if squeeze else idx.shape[0]
    b, n, _ = xyz.shape
    m, ns = idx.shape[-2], idx.shape[-1]
    (w1, b1), (w2p, b2, c2), (w3p, b3, c3) = layers
    for t, name in ((w1, "w1"), (b1, "b1"), (w2p, "w2"), (b2, "b2"), (w3p, "w3"), (b3, "b3")):
        _chk(t, name, torch.float32)
    c1 = w1.shape[0]
    if rot is not None:
        _chk(rot, "rot", torch.float32)
    out = torch.empty((b, c3, m, g), dtype=torch.float32, device=xyz.device)
    out_t = torch.empty((b, m, g, c3), dtype=torch.float32, device=xyz.device) if channels_last else None
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_group_mlp_maxpool(_ptr(xyz), _ptr(new_xyz), _ptr(rot) if rot is not None else None,
                                               _ptr(idx), g, b, n, m, ns, float(radius), _ptr(w1), _ptr(b1),
                                               _ptr(w2p), _ptr(b2), _ptr(w3p), _ptr(b3), c1, c2, c3, int(passes),
                                               _ptr(out), _ptr(out_t) if channels_last else None, _stream()),
                   "group_mlp_maxpool")
    if squeeze:
        out = out.view(b, c3, m)
        out_t = out_t.view(b, m, c3) if channels_last else None
    return (out, out_t) if channels_last else out


def group_feat_mlp_maxpool(xyz, new_xyz, feat_t, idx, layers, radius=0.0, passes=3, channels_last=False):
    """The same with point features: feat_t (B,N,C) CHANNELS-LAST, idx (B,M,ns);
    layers = [(w1_packed over [features | xyz | pad], b1, C1), (w2_packed, b2, C2), (w3_packed, b3, C3)]
    -> (B,C3,M) [, (B,M,C3)]."""
    _chk(xyz, "xyz", torch.float32)
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(feat_t, "feat_t", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, n, _ = xyz.shape
    m, ns = idx.shape[-2], idx.shape[-1]
    c = feat_t.shape[2]
    if feat_t.shape[:2] != (b, n):
        raise RuntimeError("feat_t must be (B,N,C) channels-last")
    (w1p, b1, c1), (w2p, b2, c2), (w3p, b3, c3) = layers
    for t, name in ((w1p, "w1"), (b1, "b1"), (w2p, "w2"), (b2, "b2"), (w3p, "w3"), (b3, "b3")):
        _chk(t, name, torch.float32)
    out = torch.empty((b, c3, m), dtype=torch.float32, device=xyz.device)
    out_t = torch.empty((b, m, c3), dtype=torch.float32, device=xyz.device) if channels_last else None
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_group_feat_mlp_maxpool(_ptr(xyz), _ptr(new_xyz), _ptr(feat_t), _ptr(idx), b, n, m, ns, c,
                                                    float(radius), _ptr(w1p), _ptr(b1), _ptr(w2p), _ptr(b2),
                                                    _ptr(w3p), _ptr(b3), c1, c2, c3, int(passes), _ptr(out),
                                                    _ptr(out_t) if channels_last else None, _stream()),
                   "group_feat_mlp_maxpool")
    return (out, out_t) if channels_last else out


class F16Layer:
    """One layer y = act(W x + b) prepared for the fp16x3 tensor-core kernels (csrc/mlp_f16.cu): W split hi | lo in
    fp16 after an exact power-of-two scale, packed in K-slabs; the zero-padded bias; the two numbers of the range
    bound (max row 1-norm of W, max |b|)."""

    def __init__(self, w, b, relu=True, k_pad=None, feat_channels=0, n_pad=None):
        _chk(w, "w", torch.float32)
        _chk(b, "b", torch.float32)
        cout, cin = w.shape
        self.cout, self.cin, self.relu = cout, cin, bool(relu)
        self.n_pad = (cout + 15) // 16 * 16 if n_pad is None else n_pad
        self.k_pad = (cin + 31) // 32 * 32 if k_pad is None else k_pad
        if self.n_pad > 256:
            raise RuntimeError("b200grasp: fp16x3 layers are at most 256 channels wide (split the output channels)")
        wmax = float(w.abs().max()) if w.numel() else 0.0
        self.w_exp = max(-40, min(40, 14 - math.frexp(wmax)[1])) if wmax > 0 and math.isfinite(wmax) else 0
        self.w_norm1 = float(w.abs().sum(dim=1).max()) * (1.0 + 1e-5)
        self.b_max = float(b.abs().max())
        self.bias = torch.zeros(self.n_pad, dtype=torch.float32, device=w.device)
        self.bias[:cout] = b
        nbytes = _lib.b200_mlp_f16_packed_bytes(self.n_pad, self.k_pad)
        self.packed = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
        with torch.cuda.device(w.device):
            _abi.check(_lib.b200_mlp_f16_pack(_ptr(w), cout, cin, cin, self.n_pad, self.k_pad, self.w_exp,
                                              int(feat_channels), _ptr(self.packed), _stream()), "mlp_f16_pack")

    def fill(self, dst, y=None, y_batch_stride=0):
        dst.w_packed, dst.bias = self.packed.data_ptr(), self.bias.data_ptr()
        dst.k_pad, dst.n_pad, dst.cout, dst.relu, dst.w_exp = self.k_pad, self.n_pad, self.cout, int(self.relu), self.w_exp
        dst.w_norm1, dst.b_max = self.w_norm1, self.b_max
        dst.y = y.data_ptr() if y is not None else None
        dst.y_batch_stride = int(y_batch_stride)


def f16_stack(folded, relu_last=True, first_k_pad=None, feat_channels=0, n_pads=None):
    """[(w, b)] of consecutive layers -> [F16Layer] whose K paddings chain (layer l reads layer l-1's padded width);
    n_pads: padded output widths to use instead of the next multiple of 16 (two heads of one launch share them)"""
    out = []
    for i, (w, b) in enumerate(folded):
        k_pad = first_k_pad if i == 0 else (out[-1].n_pad + 31) // 32 * 32
        out.append(F16Layer(w.contiguous(), b.contiguous(), relu=relu_last or i + 1 < len(folded), k_pad=k_pad,
                            feat_channels=feat_channels if i == 0 else 0, n_pad=None if n_pads is None else n_pads[i]))
    return out


def _layer_array(layers, ys=None, strides=None):
    arr = (_abi.MlpLayer * len(layers))()
    for i, l in enumerate(layers):
        l.fill(arr[i], None if ys is None else ys[i], 0 if strides is None else strides[i])
    return arr


def group_mlp_maxpool_f16(xyz, new_xyz, idx, first, layers, radius=0.0, rot=None, channels_last=False):
    """group_mlp_maxpool in fp16x3 arithmetic: first = (w1 (C1,3), b1 (C1), |w1| row 1-norm max, max |b1|) for the
    CUDA-core layer, layers = [F16Layer, F16Layer]."""
    _chk(xyz, "xyz", torch.float32)
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(idx, "idx", torch.int32)
    squeeze = idx.dim() == 3
    g = 1 if squeeze else idx.shape[0]
    b, n, _ = xyz.shape
    m, ns = idx.shape[-2], idx.shape[-1]
    w1, b1, w1n, b1m = first
    _chk(w1, "w1", torch.float32)
    _chk(b1, "b1", torch.float32)
    if rot is not None:
        _chk(rot, "rot", torch.float32)
    c3 = layers[-1].cout
    out = torch.empty((b, c3, m, g), dtype=torch.float32, device=xyz.device)
    out_t = torch.empty((b, m, g, c3), dtype=torch.float32, device=xyz.device) if channels_last else None
    rb = torch.empty((b * m * g, 8), dtype=torch.float32, device=xyz.device) if channels_last else None
    arr = _layer_array(layers)
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_group_mlp_maxpool_f16(_ptr(xyz), _ptr(new_xyz), _ptr(rot) if rot is not None else None,
                                                   _ptr(idx), g, b, n, m, ns, float(radius), _ptr(w1), _ptr(b1),
                                                   w1.shape[0], float(w1n), float(b1m), arr, _ptr(out),
                                                   _ptr(out_t) if channels_last else None,
                                                   _ptr(rb) if rb is not None else None, _stream()),
                   "group_mlp_maxpool_f16")
    if squeeze:
        out = out.view(b, c3, m)
        out_t = out_t.view(b, m, c3) if channels_last else None
    if channels_last and rb is not None:
        out_t.b200_row_bound = rb                      # the rider travels with THIS tensor object only (see below)
    return (out, out_t) if channels_last else out


def group_feat_mlp_maxpool_f16(xyz, new_xyz, feat_t, idx, layers, radius=0.0, channels_last=False):
    """group_feat_mlp_maxpool in fp16x3 arithmetic: layers = three F16Layer, the first packed with feat_channels = C.
    Row-bound rider: the fp16x3 grouped kernels leave, beside a channels-last output, each row's largest magnitude
    (attribute `b200_row_bound` of the returned out_t: (rows, 8) per-warp partial maxima).  When feat_t is such a tensor
    OBJECT (not a view / slice / copy of it: those do not carry the attribute) the kernel reads the 32-byte rider
    instead of scanning every gathered row for its magnitude; the result is bit-identical either way."""
    _chk(xyz, "xyz", torch.float32)
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(feat_t, "feat_t", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, n, _ = xyz.shape
    m, ns = idx.shape[-2], idx.shape[-1]
    c = feat_t.shape[2]
    if feat_t.shape[:2] != (b, n):
        raise RuntimeError("feat_t must be (B,N,C) channels-last")
    c3 = layers[-1].cout
    out = torch.empty((b, c3, m), dtype=torch.float32, device=xyz.device)
    out_t = torch.empty((b, m, c3), dtype=torch.float32, device=xyz.device) if channels_last else None
    rb = torch.empty((b * m, 8), dtype=torch.float32, device=xyz.device) if channels_last else None
    rb_in = getattr(feat_t, "b200_row_bound", None)
    if rb_in is not None and (rb_in.shape != (b * n, 8) or rb_in.device != feat_t.device):
        rb_in = None
    arr = _layer_array(layers)
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_group_feat_mlp_maxpool_f16(_ptr(xyz), _ptr(new_xyz), _ptr(feat_t),
                                                        _ptr(rb_in) if rb_in is not None else None, _ptr(idx), b, n, m,
                                                        ns, c, float(radius), arr, _ptr(out),
                                                        _ptr(out_t) if channels_last else None,
                                                        _ptr(rb) if channels_last else None, _stream()),
                   "group_feat_mlp_maxpool_f16")
    if channels_last:
        out_t.b200_row_bound = rb
    return (out, out_t) if channels_last else out


def pointwise_stack_f16(x, layers, x2=None, keep=None, out=None, out_channel_offset=0, layers2=None, out2=None,
                        out2_channel_offset=0):
    """[x (B,C1,n) | x2 (B,C2,n)] through 1..3 F16Layer in one launch -> the last layer's output (B,Cout,n); with
    keep = [bool per layer] a list of every kept layer's output.  `out` (B,Ctot,n) + out_channel_offset: write the LAST
    layer's channels there instead (layers whose output channels are split over several stacks).
    layers2: a second stack of the same padded shapes over the same input, run by the same launch (its last layer
    goes to out2 at out2_channel_offset, or to a new tensor); returns (first result, second result)."""
    _chk(x, "x", torch.float32)
    b, c1, n = x.shape
    c2 = 0
    if x2 is not None:
        _chk(x2, "x2", torch.float32)
        c2 = x2.shape[1]
    nl = len(layers)
    keep = [False] * (nl - 1) + [True] if keep is None else list(keep)
    ys, strides = [], []
    for i, l in enumerate(layers):
        if i + 1 == nl and out is not None:
            ys.append(out[:, out_channel_offset:, :])
            strides.append(out.shape[1] * n)
        elif keep[i]:
            ys.append(torch.empty((b, l.cout, n), dtype=torch.float32, device=x.device))
            strides.append(l.cout * n)
        else:
            ys.append(None)
            strides.append(0)
    heads = 1 if layers2 is None else 2
    arr = (_abi.MlpLayer * (heads * nl))()
    for i, l in enumerate(layers):
        l.fill(arr[i], None, strides[i])
        if ys[i] is not None:
            arr[i].y = ys[i].data_ptr()
    y2 = None
    if layers2 is not None:
        if len(layers2) != nl:
            raise RuntimeError("b200grasp: the two stacks of a launch must have the same depth")
        for i, l in enumerate(layers2):
            l.fill(arr[nl + i], None, 0)
        if out2 is not None:
            y2, view, stride2 = out2, out2[:, out2_channel_offset:, :], out2.shape[1] * n
        else:
            y2 = torch.empty((b, layers2[-1].cout, n), dtype=torch.float32, device=x.device)
            view, stride2 = y2, layers2[-1].cout * n
        arr[2 * nl - 1].y = view.data_ptr()
        arr[2 * nl - 1].y_batch_stride = stride2
    with torch.cuda.device(x.device):
        _abi.check(_lib.b200_pointwise_stack_f16(_ptr(x), c1, _ptr(x2) if x2 is not None else None, c2, b, n, arr,
                                                 nl, heads, _stream()), "pointwise_stack_f16")
    kept = [y for y, k in zip(ys, keep) if k]
    first = out if out is not None else (kept[0] if len(kept) == 1 else kept)
    return first if layers2 is None else (first, y2)


def gather_points_grad(grad_out, idx, n):
    """grad_out (B,C,M) f32, idx (B,M) i32 -> (B,C,n)   [sampling.cpp:45-69]"""
    _chk(grad_out, "grad_out", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, c, m = grad_out.shape
    out = torch.empty((b, c, n), dtype=torch.float32, device=grad_out.device)
    with torch.cuda.device(grad_out.device):
        _abi.check(_lib.b200_gather_points_grad(_ptr(grad_out), _ptr(idx), b, c, n, m, _ptr(out), _stream()),
                   "gather_points_grad")
    return out


# Clouds of at least this many points are queried through the uniform grid (query_grid.cu); smaller ones by
# the linear scan (query.cu).  Both give the same lists; tests flip this to compare them.
GRID_QUERY_MIN_N = 4096


class QueryGrid:
    """A cloud binned by query_grid_build: pass it as `grid=` to ball_query / cylinder_query[_multi] on the
    SAME xyz tensor to skip the binning (any cell edge gives the same lists)."""

    def __init__(self, xyz, ws, nbytes):
        self.ptr, self.shape, self.ws, self.nbytes = xyz.data_ptr(), tuple(xyz.shape), ws, nbytes

    def check(self, xyz):
        if xyz.data_ptr() != self.ptr or tuple(xyz.shape) != self.shape:
            raise ValueError("b200grasp: this QueryGrid was built from another xyz tensor")


def _grid_workspace(b, n, device):
    nbytes = _lib.b200_query_grid_workspace(b, n)
    return torch.empty((nbytes,), dtype=torch.uint8, device=device), nbytes


def query_grid_build(xyz, cell):
    """xyz (B,N,3) f32 -> QueryGrid with cell edge `cell` (grown by the library if the cloud is too wide)."""
    _chk(xyz, "xyz", torch.float32)
    b, n, _ = xyz.shape
    ws, nbytes = _grid_workspace(b, n, xyz.device)
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_query_grid_build(_ptr(xyz), b, n, float(cell), _ptr(ws), nbytes, _stream()),
                   "query_grid_build")
    return QueryGrid(xyz, ws, nbytes)


def ball_query(new_xyz, xyz, radius, nsample, grid=None):
    """new_xyz (B,M,3), xyz (B,N,3) -> (B,M,nsample) i32   [ball_query.cpp:13-37]"""
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(xyz, "xyz", torch.float32)
    b, m, _ = new_xyz.shape
    n = xyz.shape[1]
    out = torch.empty((b, m, nsample), dtype=torch.int32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        if grid is not None:
            grid.check(xyz)
            _abi.check(_lib.b200_ball_query_prebuilt(_ptr(new_xyz), _ptr(xyz), b, n, m, float(radius), int(nsample),
                                                     _ptr(out), _ptr(grid.ws), grid.nbytes, _stream()),
                       "ball_query_prebuilt")
        elif n >= GRID_QUERY_MIN_N:
            ws, nbytes = _grid_workspace(b, n, xyz.device)
            _abi.check(_lib.b200_ball_query_grid(_ptr(new_xyz), _ptr(xyz), b, n, m, float(radius), int(nsample),
                                                 _ptr(out), _ptr(ws), nbytes, _stream()), "ball_query_grid")
        else:
            _abi.check(_lib.b200_ball_query(_ptr(new_xyz), _ptr(xyz), b, n, m, float(radius), int(nsample),
                                            _ptr(out), _stream()), "ball_query")
    return out


def cylinder_query(new_xyz, xyz, rot, radius, hmin, hmax, nsample, grid=None):
    """new_xyz (B,M,3), xyz (B,N,3), rot (B,M,9) -> (B,M,nsample) i32   [cylinder_query.cpp:10-37]"""
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(xyz, "xyz", torch.float32)
    _chk(rot, "rot", torch.float32)
    b, m, _ = new_xyz.shape
    n = xyz.shape[1]
    if grid is not None or n >= GRID_QUERY_MIN_N:
        return cylinder_query_multi(new_xyz, xyz, rot, radius, hmin, [hmax], nsample, grid=grid)[0]
    out = torch.empty((b, m, nsample), dtype=torch.int32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        _abi.check(_lib.b200_cylinder_query(_ptr(new_xyz), _ptr(xyz), _ptr(rot), b, n, m, float(radius),
                                            float(hmin), float(hmax), int(nsample), _ptr(out), _stream()),
                   "cylinder_query")
    return out


def cylinder_query_multi(new_xyz, xyz, rot, radius, hmin, hmax_list, nsample, grid=None):
    """One scan for all hmax of CloudCrop (models/modules.py:127-130) -> (H,B,M,nsample) i32."""
    _chk(new_xyz, "new_xyz", torch.float32)
    _chk(xyz, "xyz", torch.float32)
    _chk(rot, "rot", torch.float32)
    b, m, _ = new_xyz.shape
    n = xyz.shape[1]
    nh = len(hmax_list)
    hm = (ctypes.c_float * max(nh, 1))(*[float(h) for h in hmax_list])
    out = torch.empty((nh, b, m, nsample), dtype=torch.int32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        if grid is not None:
            grid.check(xyz)
            _abi.check(_lib.b200_cylinder_query_multi_prebuilt(_ptr(new_xyz), _ptr(xyz), _ptr(rot), b, n, m,
                                                               float(radius), float(hmin), hm, nh, int(nsample),
                                                               _ptr(out), _ptr(grid.ws), grid.nbytes, _stream()),
                       "cylinder_query_multi_prebuilt")
        elif n >= GRID_QUERY_MIN_N:
            ws, nbytes = _grid_workspace(b, n, xyz.device)
            _abi.check(_lib.b200_cylinder_query_multi_grid(_ptr(new_xyz), _ptr(xyz), _ptr(rot), b, n, m,
                                                           float(radius), float(hmin), hm, nh, int(nsample),
                                                           _ptr(out), _ptr(ws), nbytes, _stream()),
                       "cylinder_query_multi_grid")
        else:
            _abi.check(_lib.b200_cylinder_query_multi(_ptr(new_xyz), _ptr(xyz), _ptr(rot), b, n, m, float(radius),
                                                      float(hmin), hm, nh, int(nsample), _ptr(out), _stream()),
                       "cylinder_query_multi")
    return out


def group_points(points, idx):
    """points (B,C,N) f32, idx (B,M,ns) i32 -> (B,C,M,ns)   [group_points.cpp:17-40]"""
    _chk(points, "points", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, c, n = points.shape
    _, m, ns = idx.shape
    out = torch.empty((b, c, m, ns), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        _abi.check(_lib.b200_group_points(_ptr(points), _ptr(idx), b, c, n, m, ns, _ptr(out), _stream()),
                   "group_points")
    return out


def group_points_grad(grad_out, idx, n):
    """grad_out (B,C,M,ns) f32, idx (B,M,ns) i32 -> (B,C,n)   [group_points.cpp:42-65]"""
    _chk(grad_out, "grad_out", torch.float32)
    _chk(idx, "idx", torch.int32)
    b, c, m, ns = grad_out.shape
    out = torch.empty((b, c, n), dtype=torch.float32, device=grad_out.device)
    with torch.cuda.device(grad_out.device):
        _abi.check(_lib.b200_group_points_grad(_ptr(grad_out), _ptr(idx), b, c, n, m, ns, _ptr(out), _stream()),
                   "group_points_grad")
    return out


def three_nn(unknowns, knows):
    """unknowns (B,n,3), knows (B,m,3) -> (dist2 (B,n,3) f32, idx (B,n,3) i32)   [interpolate.cpp:19-45]"""
    _chk(unknowns, "unknowns", torch.float32)
    _chk(knows, "knows", torch.float32)
    b, n, _ = unknowns.shape
    m = knows.shape[1]
    dist2 = torch.empty((b, n, 3), dtype=torch.float32, device=unknowns.device)
    idx = torch.empty((b, n, 3), dtype=torch.int32, device=unknowns.device)
    with torch.cuda.device(unknowns.device):
        _abi.check(_lib.b200_three_nn(_ptr(unknowns), _ptr(knows), b, n, m, _ptr(dist2), _ptr(idx), _stream()),
                   "three_nn")
    return dist2, idx


def pointwise_conv(x, w, b, relu=True, x2=None):
    """1x1 convolution over point columns: x (B,C1,n) [concatenated with x2 (B,C2,n) along channels], w (Cout,C1+C2)
    -- or (Cout, C1+C2 rounded up to a multiple of 4) with ZERO padding columns, which takes the 16-byte copy path
    (pad_weight) --, b (Cout) -> [relu](w x + b) (B,Cout,n); bias / folded BatchNorm and ReLU in the epilogue"""
    _chk(x, "x", torch.float32)
    _chk(w, "w", torch.float32)
    _chk(b, "b", torch.float32)
    bsz, c1, n = x.shape
    c2 = 0
    if x2 is not None:
        _chk(x2, "x2", torch.float32)
        if x2.shape[0] != bsz or x2.shape[2] != n:
            raise ValueError("b200grasp: x and x2 must agree in batch and columns")
        c2 = x2.shape[1]
    cout, ldw = w.shape
    if not c1 + c2 <= ldw < c1 + c2 + 4 or b.numel() != cout:
        raise ValueError("b200grasp: weight (%d,%d) does not match %d input channels" % (cout, ldw, c1 + c2))
    y = torch.empty((bsz, cout, n), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _abi.check(_lib.b200_pointwise_conv(_ptr(w), ldw, _ptr(b), _ptr(x), c1, _ptr(x2) if x2 is not None else None, c2,
                                            bsz, n, cout, 1 if relu else 0, _ptr(y), _stream()), "pointwise_conv")
    return y


def pad_weight(w):
    """(Cout,Cin) -> (Cout, Cin rounded up to a multiple of 4), zero-padded, contiguous (see pointwise_conv)"""
    pad = (-w.shape[1]) % 4
    return torch.nn.functional.pad(w, (0, pad)).contiguous() if pad else w.contiguous()


def three_nn_weights(dist2):
    """dist2 (B,n,3) from three_nn -> inverse-distance weights (B,n,3) = (1/(sqrt(dist2)+1e-8)) / sum
    [pointnet2_modules.py:394-397 after the sqrt of pointnet2_utils.py:140], one launch"""
    _chk(dist2, "dist2", torch.float32)
    weight = torch.empty_like(dist2)
    with torch.cuda.device(dist2.device):
        _abi.check(_lib.b200_three_nn_weights(_ptr(dist2), dist2.numel() // 3, None, _ptr(weight), _stream()),
                   "three_nn_weights")
    return weight


def view_select(features, views, frames):
    """features (B,2+V,S) f32 (ApproachNet conv3 output), views (V,3), frames (V,9) ->
    (top_view_inds (B,S) i64, top_view_score (B,S), vp_xyz (B,S,3), vp_rot (B,S,3,3))   [models/modules.py:60-73]"""
    _chk(features, "features", torch.float32)
    _chk(views, "views", torch.float32)
    _chk(frames, "frames", torch.float32)
    b, c, s = features.shape
    v = c - 2
    if tuple(views.shape) != (v, 3) or frames.numel() != v * 9:
        raise ValueError("b200grasp: view tables do not match the %d view scores" % v)
    dev = features.device
    inds = torch.empty((b, s), dtype=torch.int64, device=dev)
    score = torch.empty((b, s), dtype=torch.float32, device=dev)
    xyz = torch.empty((b, s, 3), dtype=torch.float32, device=dev)
    rot = torch.empty((b, s, 3, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _abi.check(_lib.b200_view_select(_ptr(features), _ptr(views), _ptr(frames), b, s, v, _ptr(inds), _ptr(score),
                                         _ptr(xyz), _ptr(rot), _stream()), "view_select")
    return inds, score, xyz, rot


def three_interpolate(points, idx, weight):
    """points (B,c,m), idx (B,n,3) i32, weight (B,n,3) -> (B,c,n)   [interpolate.cpp:47-75]"""
    _chk(points, "points", torch.float32)
    _chk(idx, "idx", torch.int32)
    _chk(weight, "weight", torch.float32)
    b, c, m = points.shape
    n = idx.shape[1]
    out = torch.empty((b, c, n), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        _abi.check(_lib.b200_three_interpolate(_ptr(points), _ptr(idx), _ptr(weight), b, c, m, n, _ptr(out),
                                               _stream()), "three_interpolate")
    return out


def three_interpolate_grad(grad_out, idx, weight, m):
    """grad_out (B,c,n), idx, weight (B,n,3) -> (B,c,m)   [interpolate.cpp:76-104]"""
    _chk(grad_out, "grad_out", torch.float32)
    _chk(idx, "idx", torch.int32)
    _chk(weight, "weight", torch.float32)
    b, c, n = grad_out.shape
    out = torch.empty((b, c, m), dtype=torch.float32, device=grad_out.device)
    with torch.cuda.device(grad_out.device):
        _abi.check(_lib.b200_three_interpolate_grad(_ptr(grad_out), _ptr(idx), _ptr(weight), b, c, n, m,
                                                    _ptr(out), _stream()), "three_interpolate_grad")
    return out


def knn(ref, query, idx):
    """ref (B,D,Nr), query (B,D,Nq), idx (B,k,Nq) i64 OUT, caller-allocated -> 1   [knn.h:12-65]"""
    _chk(ref, "ref", torch.float32)
    _chk(query, "query", torch.float32)
    _chk(idx, "idx", torch.int64)
    b, d, nr = ref.shape
    nq = query.shape[2]
    k = idx.shape[1]
    with torch.cuda.device(ref.device):
        _abi.check(_lib.b200_knn(_ptr(ref), _ptr(query), b, d, nr, nq, k, _ptr(idx), _stream()), "knn")
    return 1


def collision_counts(scene, T, R, thr, n_dev=None):
    """scene (N,3) f64, T (M,3), R (M,9), thr (M,10) f64 -> counts (M,6) i32
    [utils/collision_detector.py:75-100,119].  n_dev: optional (1,) i32 device tensor holding the
    number of valid scene rows (as written by voxel_down_sample), so no host round trip is needed."""
    for t, nm in ((scene, "scene"), (T, "T"), (R, "R"), (thr, "thr")):
        _chk(t, nm, torch.float64)
    n, m = scene.shape[0], T.shape[0]
    counts = torch.empty((m, 6), dtype=torch.int32, device=scene.device)
    with torch.cuda.device(scene.device):
        if n_dev is None:
            _abi.check(_lib.b200_collision_counts(_ptr(scene), n, _ptr(T), _ptr(R), _ptr(thr), m, _ptr(counts),
                                                  _stream()), "collision_counts")
        else:
            _chk(n_dev, "n_dev", torch.int32)
            _abi.check(_lib.b200_collision_counts_dn(_ptr(scene), _ptr(n_dev), n, _ptr(T), _ptr(R), _ptr(thr), m,
                                                     _ptr(counts), _stream()), "collision_counts_dn")
    return counts


def collision_counts_packed(scene, packed, n_dev=None):
    """scene (N,3) f64, packed (M,22) f64 = T | R | thr per grasp -> counts (M,6) i32 (see collision_counts)."""
    _chk(scene, "scene", torch.float64)
    _chk(packed, "packed", torch.float64)
    if packed.dim() != 2 or packed.shape[1] != 22:
        raise ValueError("b200grasp: packed must be (M,22)")
    n, m = scene.shape[0], packed.shape[0]
    counts = torch.empty((m, 6), dtype=torch.int32, device=scene.device)
    if n_dev is not None:
        _chk(n_dev, "n_dev", torch.int32)
    with torch.cuda.device(scene.device):
        _abi.check(_lib.b200_collision_counts_packed(_ptr(scene), _ptr(n_dev) if n_dev is not None else None, n,
                                                     _ptr(packed), m, _ptr(counts), _stream()),
                   "collision_counts_packed")
    return counts


def pred_decode(objectness, score, angle_cls, width, tolerance, xyz, view_xyz):
    """Head outputs of a batch -> (rows (B,Ns,17) f32, nrows (B,) i32), both on the device
    [models/graspnet.py:77-134].  objectness (B,2,Ns); score/angle_cls/width/tolerance (B,A,Ns,D); xyz, view (B,Ns,3)."""
    for t, nm in ((objectness, "objectness"), (score, "score"), (angle_cls, "angle_cls"), (width, "width"),
                  (tolerance, "tolerance"), (xyz, "xyz"), (view_xyz, "view_xyz")):
        _chk(t, nm, torch.float32)
    b, na, ns, nd = score.shape
    if tuple(objectness.shape) != (b, 2, ns) or tuple(xyz.shape) != (b, ns, 3) or tuple(view_xyz.shape) != (b, ns, 3) \
            or any(tuple(t.shape) != (b, na, ns, nd) for t in (angle_cls, width, tolerance)):
        raise ValueError("b200grasp: pred_decode shapes do not agree")
    rows = torch.empty((b, ns, 17), dtype=torch.float32, device=score.device)
    nrows = torch.empty((b,), dtype=torch.int32, device=score.device)
    with torch.cuda.device(score.device):
        _abi.check(_lib.b200_pred_decode(_ptr(objectness), _ptr(score), _ptr(angle_cls), _ptr(width), _ptr(tolerance),
                                         _ptr(xyz), _ptr(view_xyz), b, ns, na, nd, _ptr(rows), _ptr(nrows), _stream()),
                   "pred_decode")
    return rows, nrows


def collision_mask_device(scene, n_dev, grasps, m_dev, voxel_size, approach_dist, collision_thresh, out=None):
    """scene (N,3) f64 (+ n_dev (1,) i32 or None), grasps (Mmax,17) f32 of which m_dev (1,) i32 rows are valid ->
    mask (Mmax,) u8 on the device, 1 = collides   [collision_detector.py:37-112, default return].
    `out`: a (Mmax,) u8 buffer to write into (only the first m_dev entries are written)."""
    _chk(scene, "scene", torch.float64)
    _chk(grasps, "grasps", torch.float32)
    _chk(m_dev, "m_dev", torch.int32)
    if n_dev is not None:
        _chk(n_dev, "n_dev", torch.int32)
    m = grasps.shape[0]
    if out is None:
        mask = torch.zeros((m,), dtype=torch.uint8, device=grasps.device)
    else:
        _chk(out, "out", torch.uint8)
        if out.numel() != m:
            raise ValueError("b200grasp: out must hold one byte per grasp row")
        mask = out
    nbytes = _lib.b200_collision_mask_workspace(m)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=grasps.device)
    with torch.cuda.device(grasps.device):
        _abi.check(_lib.b200_collision_mask_device(_ptr(scene), _ptr(n_dev) if n_dev is not None else None,
                                                   scene.shape[0], _ptr(grasps), _ptr(m_dev), m, float(voxel_size),
                                                   float(approach_dist), float(collision_thresh), _ptr(mask), _ptr(ws),
                                                   nbytes, _stream()), "collision_mask_device")
    return mask


def compact_rows(rows, drop, m_dev):
    """rows (Mmax,C) f32, drop (Mmax,) u8, m_dev (1,) i32 -> (rows[~drop] in order, padded to Mmax; count (1,) i32)"""
    _chk(rows, "rows", torch.float32)
    _chk(drop, "drop", torch.uint8)
    _chk(m_dev, "m_dev", torch.int32)
    m, c = rows.shape
    out = torch.empty_like(rows)
    nout = torch.empty((1,), dtype=torch.int32, device=rows.device)
    with torch.cuda.device(rows.device):
        _abi.check(_lib.b200_compact_rows(_ptr(rows), c, _ptr(drop), _ptr(m_dev), m, _ptr(out), _ptr(nout), _stream()),
                   "compact_rows")
    return out, nout


def grasp_nms_device(grasps, m_dev, translation_thresh, cos_rotation_thresh, out=None, out_count=None):
    """grasps (Mmax,17) f32 with m_dev (1,) i32 valid rows -> (kept rows (Mmax,17) in score order, nkeep (1,) i32);
    `out` / `out_count`: buffers of those shapes to write into"""
    _chk(grasps, "grasps", torch.float32)
    _chk(m_dev, "m_dev", torch.int32)
    m = grasps.shape[0]
    keep = torch.empty((max(m, 1),), dtype=torch.int32, device=grasps.device)
    if out_count is None:
        nkeep = torch.empty((1,), dtype=torch.int32, device=grasps.device)
    else:
        _chk(out_count, "out_count", torch.int32)
        nkeep = out_count
    if out is None:
        out = torch.empty_like(grasps)
    else:
        _chk(out, "out", torch.float32)
        if tuple(out.shape) != tuple(grasps.shape):
            raise ValueError("b200grasp: out must have the shape of the grasp rows")
    nbytes = _lib.b200_grasp_nms_workspace(m)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=grasps.device)
    with torch.cuda.device(grasps.device):
        _abi.check(_lib.b200_grasp_nms_device(_ptr(grasps), _ptr(m_dev), m, float(translation_thresh),
                                              float(cos_rotation_thresh), _ptr(keep), _ptr(nkeep), _ptr(ws), nbytes,
                                              _stream()), "grasp_nms_device")
        _abi.check(_lib.b200_gather_rows(_ptr(grasps), grasps.shape[1], _ptr(keep), _ptr(nkeep), m, _ptr(out),
                                         _stream()), "gather_rows")
    return out, nkeep


POST_CHAIN_MAX_M = 2048                  # B200_POST_CHAIN_MAX_M (include/b200grasp.h)


def grasp_post_chain(scene, n_dev, grasps, m_dev, voxel_size, approach_dist, collision_thresh, translation_thresh,
                     cos_rotation_thresh, out_mask=None, out=None, out_count=None, out_nrows=None):
    """collision_mask_device + compact_rows + grasp_nms_device in one call and five graph nodes (csrc/post_chain.cu;
    same arithmetic, same results): scene (N,3) f64 (+ n_dev), grasps (Mmax <= POST_CHAIN_MAX_M, 17) f32 with m_dev
    valid rows -> (kept rows (Mmax,17) in score order, nkept (1,) i32, collision mask (Mmax,) u8).
    `out_mask` / `out` / `out_count`: buffers to write into; `out_nrows`: (1,) i32 that receives the valid row count."""
    _chk(scene, "scene", torch.float64)
    _chk(grasps, "grasps", torch.float32)
    _chk(m_dev, "m_dev", torch.int32)
    if n_dev is not None:
        _chk(n_dev, "n_dev", torch.int32)
    m = grasps.shape[0]
    if m > POST_CHAIN_MAX_M:
        raise ValueError("b200grasp: grasp_post_chain handles at most %d rows (use the separate calls)" % POST_CHAIN_MAX_M)
    dev = grasps.device
    mask = torch.zeros((m,), dtype=torch.uint8, device=dev) if out_mask is None else out_mask
    kept = torch.empty_like(grasps) if out is None else out
    nkept = torch.empty((1,), dtype=torch.int32, device=dev) if out_count is None else out_count
    _chk(mask, "out_mask", torch.uint8)
    _chk(kept, "out", torch.float32)
    _chk(nkept, "out_count", torch.int32)
    if out_nrows is not None:
        _chk(out_nrows, "out_nrows", torch.int32)
    if mask.numel() != m or tuple(kept.shape) != tuple(grasps.shape):
        raise ValueError("b200grasp: output buffers must match the grasp rows")
    nbytes = _lib.b200_grasp_post_chain_workspace(m)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _abi.check(_lib.b200_grasp_post_chain(_ptr(scene), _ptr(n_dev) if n_dev is not None else None, scene.shape[0],
                                              _ptr(grasps), _ptr(m_dev), m, float(voxel_size), float(approach_dist),
                                              float(collision_thresh), float(translation_thresh),
                                              float(cos_rotation_thresh), _ptr(mask), _ptr(kept), _ptr(nkept),
                                              _ptr(out_nrows) if out_nrows is not None else None, _ptr(ws), nbytes,
                                              _stream()), "grasp_post_chain")
    return kept, nkept, mask


def voxel_down_sample(points, voxel_size, return_bad=False):
    """points (N,3) f64 -> (out (N,3) f64 of which the first nvox rows are valid, nvox (1,) i32), both on
    the device   [utils/collision_detector.py:32-35 -> open3d voxel_down_sample, restated].
    return_bad: also nbad (1,) i32 = points left out (not finite, or beyond 2^21 voxels per axis)."""
    _chk(points, "points", torch.float64)
    n = points.shape[0]
    out = torch.empty((max(n, 1), 3), dtype=torch.float64, device=points.device)
    counts = torch.empty((2,), dtype=torch.int32, device=points.device)
    nvox, nbad = counts[0:1], counts[1:2]
    nbytes = _lib.b200_voxel_down_sample_workspace(n)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=points.device)
    with torch.cuda.device(points.device):
        _abi.check(_lib.b200_voxel_down_sample_checked(_ptr(points), n, float(voxel_size), _ptr(out), _ptr(nvox),
                                                       _ptr(nbad), _ptr(ws), nbytes, _stream()), "voxel_down_sample")
    return (out, nvox, nbad) if return_bad else (out, nvox)


def grasp_nms(grasps, translation_thresh, cos_rotation_thresh):
    """grasps (M,17) f32 -> (keep (M,) i32 row numbers in score order, nkeep (1,) i32), both on device."""
    _chk(grasps, "grasps", torch.float32)
    m = grasps.shape[0]
    keep = torch.empty((max(m, 1),), dtype=torch.int32, device=grasps.device)
    nkeep = torch.empty((1,), dtype=torch.int32, device=grasps.device)
    nbytes = _lib.b200_grasp_nms_workspace(m)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=grasps.device)
    with torch.cuda.device(grasps.device):
        _abi.check(_lib.b200_grasp_nms(_ptr(grasps), m, float(translation_thresh), float(cos_rotation_thresh),
                                       _ptr(keep), _ptr(nkeep), _ptr(ws), nbytes, _stream()), "grasp_nms")
    return keep, nkeep


def enable_nvtx(on=True):
    """Wrap every public op of this module in an NVTX range named after it (per-op ranges for
    nsys / ncu --nvtx).  Also switched on by the environment variable B200GRASP_NVTX=1 at import."""
    import functools
    import inspect
    import sys
    mod = sys.modules[__name__]
    for name, fn in list(vars(mod).items()):
        if name.startswith("_") or name == "enable_nvtx" or not inspect.isfunction(fn) or fn.__module__ != __name__:
            continue
        base = getattr(fn, "__wrapped_op__", fn)
        if not on:
            setattr(mod, name, base)
            continue

        def make(f, label):
            @functools.wraps(f)
            def wrapped(*a, **k):
                torch.cuda.nvtx.range_push(label)
                try:
                    return f(*a, **k)
                finally:
                    torch.cuda.nvtx.range_pop()
            wrapped.__wrapped_op__ = f
            return wrapped
        setattr(mod, name, make(base, "b200grasp." + name))


import os as _os
if _os.environ.get("B200GRASP_NVTX") == "1":
    enable_nvtx(True)
