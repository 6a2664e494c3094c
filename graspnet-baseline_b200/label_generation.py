"""Grasp labels for the seeds of a training batch (reference: utils/label_generation.py).

    process_grasp_labels(end_points)            :18-122   nearest annotated grasp point per seed (knn, k = 1) and its labels
    match_grasp_view_and_label(end_points)      :125-149  the slices of the seed's predicted top view

Same call signatures, same end_points keys in and out.  Two implementations with the same results:
  * composed (fused=False): the reference's own sequence -- index_select over views, cat over objects, index_select
    over the knn result, the label transform on the (B, Ns, V, A, D) tensor, torch.gather of the top view -- on this
    package's knn.  ~300 MB of gathered tensors per scene.
  * fused (default): csrc/labels.cu reads the rows it needs straight from the per-object arrays (one streaming pass
    for the view labels and u_max, one small gather for the matched slices); the (B, Ns, V, A, D[, 3]) tensors are never
    built.  Keys that only exist as those big tensors between the two calls ('batch_grasp_view', and the unmatched
    'batch_grasp_label' / '_offset' / '_tolerance' / '_view_rot') are not produced; everything the loss
    (models/loss.py) reads is.
"""
import ctypes

import numpy as np
import torch

from . import _abi, ops, synth
from .knn_modules import knn

GRASP_MAX_WIDTH = 0.1            # utils/loss_utils.py:15
FUSED = True


def _template(num_view, device):
    """(views (V,3), frames (V,3,3)): the template approach directions (generate_grasp_views, loss_utils.py:41-66) and
    the frame of each (-view, angle 0) (batch_viewpoint_params_to_matrix, :68-96)"""
    from .graspnet import view_frames
    views = torch.from_numpy(synth.fibonacci_views(num_view).astype(np.float32)).to(device)
    frames = view_frames(-views, torch.zeros(num_view, dtype=views.dtype, device=device))
    return views, frames.contiguous()


def _rigid(points, pose):
    """rows of `points` (n,3) through a 3x4 pose, with the reference's expression (transform_point_cloud '3x4')"""
    ones = points.new_ones(points.size(0), 1)
    return torch.matmul(pose, torch.cat([points, ones], dim=1).T).T[:, :3]


def _nearest(ref_pts, query_pts):
    """index (0-based, int64) of the nearest row of ref_pts (n,3) for every row of query_pts (m,3): the drop-in knn"""
    ref = ref_pts.transpose(0, 1).contiguous().unsqueeze(0)
    qry = query_pts.transpose(0, 1).contiguous().unsqueeze(0)
    return knn(ref, qry, k=1).reshape(-1) - 1


def _object_views(pose, views):
    """template views in the scene frame and, for every template view, the nearest of them (label_generation.py:52-60)"""
    views_trans = torch.matmul(pose[:3, :3], views.T).T
    return views_trans, _nearest(views_trans, views)


def _process_composed(end_points):
    seeds = end_points['fp2_xyz']
    B, Ns, _ = seeds.size()
    out = {k: [] for k in ("point", "view", "rot", "label", "offset", "tol")}
    V = A = D = None
    for i in range(B):
        merged = {k: [] for k in out}
        for o, pose in enumerate(end_points['object_poses_list'][i]):
            pts = end_points['grasp_points_list'][i][o]
            labels = end_points['grasp_labels_list'][i][o]
            offsets = end_points['grasp_offsets_list'][i][o]
            tol = end_points['grasp_tolerance_list'][i][o]
            n, V, A, D = labels.size()
            views, frames = _template(V, pose.device)
            views_trans, vi = _object_views(pose, views)
            merged["point"].append(_rigid(pts, pose))
            merged["view"].append(views_trans[vi].unsqueeze(0).expand(n, -1, -1))
            merged["rot"].append(torch.matmul(pose[:3, :3], frames)[vi].unsqueeze(0).expand(n, -1, -1, -1))
            merged["label"].append(labels[:, vi])
            merged["offset"].append(offsets[:, vi])
            merged["tol"].append(tol[:, vi])
        merged = {k: torch.cat(v, dim=0) for k, v in merged.items()}
        nn = _nearest(merged["point"], seeds[i])
        for k in out:
            out[k].append(merged[k][nn])
    out = {k: torch.stack(v, 0) for k, v in out.items()}
    labels = out["label"]
    mask = (labels > 0) & (out["offset"][..., 2] <= GRASP_MAX_WIDTH)
    u_max = labels.max()
    labels[mask] = torch.log(u_max / labels[mask])
    labels[~mask] = 0
    end_points['batch_grasp_point'] = out["point"]
    end_points['batch_grasp_view'] = out["view"]
    end_points['batch_grasp_view_rot'] = out["rot"]
    end_points['batch_grasp_label'] = labels
    end_points['batch_grasp_offset'] = out["offset"]
    end_points['batch_grasp_tolerance'] = out["tol"]
    end_points['batch_grasp_view_label'] = labels.view(B, Ns, V, A * D).max(dim=-1)[0].float()
    return end_points


_TEMPLATES = {}


def _template_cached(num_view, device):
    key = (num_view, str(device))
    if key not in _TEMPLATES:
        _TEMPLATES[key] = _template(num_view, device)
    return _TEMPLATES[key]


class _LabelTables:
    """Device tables of one batch for csrc/labels.cu: per object {label / offset / tolerance pointers, view map, pose
    rotation}, per merged grasp point {object, row within the object}, per scene {first object, first point}.  Built
    with a handful of batched launches for all objects of all scenes and no device -> host copy."""

    def __init__(self, end_points, device):
        poses_list = end_points['object_poses_list']
        flat = lambda key: [t for scene in end_points[key] for t in scene]
        poses, pts = flat('object_poses_list'), flat('grasp_points_list')
        labels, offsets, tol = flat('grasp_labels_list'), flat('grasp_offsets_list'), flat('grasp_tolerance_list')
        n_obj = len(poses)
        for t, name in [(x, "grasp_labels") for x in labels] + [(x, "grasp_offsets") for x in offsets] + \
                       [(x, "grasp_tolerance") for x in tol]:
            ops._chk(t, name, torch.float32)
        _, self.V, self.A, self.D = labels[0].size()
        views, self.frames = _template_cached(self.V, device)
        pose = torch.stack(poses, 0)                                     # (O, 3, 4)
        self.rot = pose[:, :, :3].contiguous()                           # (O, 3, 3)
        # for every template view of the scene frame, the nearest rotated template view (label_generation.py:52-60)
        views_trans = torch.matmul(self.rot, views.T)                    # (O, 3, V) = (R views^T)
        vi = knn(views_trans.contiguous(), views.T.unsqueeze(0).expand(n_obj, -1, -1).contiguous(), k=1) - 1
        self.view_inds = vi.reshape(n_obj, self.V).to(torch.int32).contiguous()
        # all grasp points in their scene's frame
        counts = [int(p.size(0)) for p in pts]
        oid = torch.repeat_interleave(torch.arange(n_obj, device=device), torch.tensor(counts, device=device),
                                      output_size=sum(counts))
        allp = torch.cat(pts, 0)
        merged = torch.einsum('nij,nj->ni', self.rot[oid], allp) + pose[oid, :, 3]
        host = (ctypes.c_uint8 * _abi.lib.b200_label_object_table_bytes(n_obj))()
        obj_first, point_first, point_obj, point_local, self.merged = [], [], [], [], []
        oi = npts = 0
        for poses_b in poses_list:
            obj_first.append(oi)
            point_first.append(npts)
            n_scene = 0
            for o in range(len(poses_b)):
                _abi.check(_abi.lib.b200_label_object_table_set(
                    host, oi, ops._ptr(labels[oi]), ops._ptr(offsets[oi]), ops._ptr(tol[oi]),
                    ctypes.c_void_p(self.view_inds.data_ptr() + 4 * self.V * oi),
                    ctypes.c_void_p(self.rot.data_ptr() + 36 * oi)), "label_object_table_set")
                point_obj.append(np.full(counts[oi], o, np.int32))
                point_local.append(np.arange(counts[oi], dtype=np.int32))
                n_scene += counts[oi]
                oi += 1
            self.merged.append(merged[npts: npts + n_scene])            # per scene (Np', 3), scene frame
            npts += n_scene
        self.keep = [labels, offsets, tol]                               # what the pointer table refers to
        i32 = lambda v: torch.from_numpy(np.asarray(v, dtype=np.int32)).to(device, non_blocking=True)
        self.obj_first, self.point_first = i32(obj_first), i32(point_first)
        self.point_obj, self.point_local = i32(np.concatenate(point_obj)), i32(np.concatenate(point_local))
        self.objs = torch.frombuffer(host, dtype=torch.uint8).clone().to(device, non_blocking=True)
        self.umax_key = torch.zeros(1, dtype=torch.int32, device=device)
        self.nn = None


def _process_fused(end_points):
    seeds = end_points['fp2_xyz']
    B, Ns, _ = seeds.size()
    dev = seeds.device
    with torch.no_grad():
        tb = _LabelTables(end_points, dev)
        if len({int(m.size(0)) for m in tb.merged}) == 1:               # equal point counts: one batched search
            ref = torch.stack(tb.merged, 0).transpose(1, 2).contiguous()
            tb.nn = (knn(ref, seeds.transpose(1, 2).contiguous(), k=1).reshape(B, Ns) - 1).contiguous()
        else:
            tb.nn = torch.stack([_nearest(tb.merged[i], seeds[i]) for i in range(B)], 0).contiguous()   # (B, Ns) int64
        V, AD = tb.V, tb.A * tb.D
        lmin = torch.empty((B, Ns, V), dtype=torch.float32, device=dev)
        scores = torch.empty((B, Ns, V), dtype=torch.float32, device=dev)
        umax = torch.empty(1, dtype=torch.float32, device=dev)
        lib, st = _abi.lib, ops._stream()
        timing = end_points.get('_time_label_kernels')
        if timing:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            ev[0].record()
        with torch.cuda.device(dev):
            _abi.check(lib.b200_label_view_stats(ops._ptr(tb.objs), ops._ptr(tb.obj_first), ops._ptr(tb.point_obj),
                                                 ops._ptr(tb.point_local), ops._ptr(tb.point_first), ops._ptr(tb.nn), B, Ns,
                                                 V, AD, GRASP_MAX_WIDTH, ops._ptr(lmin), ops._ptr(tb.umax_key), st),
                       "label_view_stats")
            _abi.check(lib.b200_label_view_scores(ops._ptr(lmin), ops._ptr(tb.umax_key), lmin.numel(), ops._ptr(scores),
                                                  ops._ptr(umax), st), "label_view_scores")
        if timing:                                        # bench.py: the streaming pass alone (it synchronises)
            ev[1].record()
            ev[1].synchronize()
            end_points['_label_kernel_ms'] = ev[0].elapsed_time(ev[1])
            end_points['_label_kernel_bytes'] = B * Ns * V * (AD * 16 + 8)     # label + offset rows read, lmin + score written
        end_points['batch_grasp_point'] = torch.stack([tb.merged[i][tb.nn[i]] for i in range(B)], 0)
        end_points['batch_grasp_point'] = end_points['batch_grasp_point'].contiguous()
        end_points['batch_grasp_view_label'] = scores
        end_points['batch_grasp_label_max'] = umax
        end_points['_label_tables'] = tb
    return end_points


def process_grasp_labels(end_points, fused=None):
    """end_points with 'fp2_xyz' (B,Ns,3), 'object_poses_list', 'grasp_points_list', 'grasp_labels_list',
    'grasp_offsets_list', 'grasp_tolerance_list' (graspnet_dataset.py:238-246) -> + 'batch_grasp_point' (B,Ns,3),
    'batch_grasp_view_label' (B,Ns,V) [and, composed, the (B,Ns,V,...) tensors]"""
    if (FUSED if fused is None else fused) and end_points['fp2_xyz'].is_cuda:
        return _process_fused(end_points)
    return _process_composed(end_points)


def match_grasp_view_and_label(end_points):
    """-> (top_template_views_rot (B,Ns,3,3), labels (B,Ns,A,D), offsets (B,Ns,A,D,3), tolerance (B,Ns,A,D), end_points)
    for each seed's predicted view 'grasp_top_view_inds' (B,Ns)"""
    top = end_points['grasp_top_view_inds']
    tb = end_points.get('_label_tables')
    if tb is None:                                       # composed: torch.gather over the view axis
        rot, labels = end_points['batch_grasp_view_rot'], end_points['batch_grasp_label']
        offsets, tol = end_points['batch_grasp_offset'], end_points['batch_grasp_tolerance']
        B, Ns, V, A, D = labels.size()
        pick = lambda t, tail: torch.gather(t, 2, top.view(B, Ns, 1, *([1] * len(tail))).expand(-1, -1, -1, *tail)).squeeze(2)
        rot, labels, offsets, tol = pick(rot, (3, 3)), pick(labels, (A, D)), pick(offsets, (A, D, 3)), pick(tol, (A, D))
    else:
        B, Ns = top.shape
        dev = top.device
        A, D, V = tb.A, tb.D, tb.V
        with torch.no_grad():
            labels = torch.empty((B, Ns, A, D), dtype=torch.float32, device=dev)
            offsets = torch.empty((B, Ns, A, D, 3), dtype=torch.float32, device=dev)
            tol = torch.empty((B, Ns, A, D), dtype=torch.float32, device=dev)
            rot = torch.empty((B, Ns, 3, 3), dtype=torch.float32, device=dev)
            top64 = top.to(torch.int64).contiguous()
            with torch.cuda.device(dev):
                _abi.check(_abi.lib.b200_label_top_gather(
                    ops._ptr(tb.objs), ops._ptr(tb.obj_first), ops._ptr(tb.point_obj), ops._ptr(tb.point_local),
                    ops._ptr(tb.point_first), ops._ptr(tb.nn), ops._ptr(top64), ops._ptr(tb.frames), ops._ptr(tb.umax_key),
                    B, Ns, V, A * D, GRASP_MAX_WIDTH, ops._ptr(labels), ops._ptr(offsets), ops._ptr(tol), ops._ptr(rot),
                    ops._stream()), "label_top_gather")
    end_points['batch_grasp_view_rot'] = rot
    end_points['batch_grasp_label'] = labels
    end_points['batch_grasp_offset'] = offsets
    end_points['batch_grasp_tolerance'] = tol
    return rot, labels, offsets, tol, end_points
