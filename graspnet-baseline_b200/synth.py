"""Seeded synthetic workloads of the shapes BASELINE.json names:
tabletop clouds, approach-view rotations and 17-column grasp rows.  numpy only;
used by bench.py, __graft_entry__.smoke() and the tests."""
import numpy as np


def tabletop_cloud(n, seed=0, extent=0.3, dtype=np.float32):
    """Plane z=0.6 +- 1 mm over [-extent,extent]^2 (60 %) + 8 boxes/spheres of 5-10 cm (40 %),
    camera frame, metres; shuffled so index order carries no spatial order."""
    rng = np.random.default_rng(seed)
    npl = n * 6 // 10
    plane = np.stack([rng.uniform(-extent, extent, npl), rng.uniform(-extent, extent, npl),
                      0.6 + rng.normal(0, 0.001, npl)], 1)
    rest = n - npl
    parts = []
    for o in range(8):
        k = rest // 8 + (1 if o < rest % 8 else 0)
        c = np.array([rng.uniform(-0.7, 0.7) * extent, rng.uniform(-0.7, 0.7) * extent, 0.0])
        s = rng.uniform(0.025, 0.05)
        if o % 2 == 0:   # box surface-ish: random points in the box, z above the table
            q = rng.uniform(-s, s, (k, 3))
            q[:, 2] = 0.6 - np.abs(q[:, 2]) - 0.001
        else:            # upper hemisphere of a sphere resting on the table
            v = rng.normal(size=(k, 3))
            v /= np.linalg.norm(v, axis=1, keepdims=True)
            q = v * s
            q[:, 2] = 0.6 - s - np.abs(q[:, 2])
        q[:, :2] += c[:2]
        parts.append(q)
    pts = np.concatenate([plane] + parts)[:n]
    rng.shuffle(pts)
    return np.ascontiguousarray(pts, dtype=dtype)


def adversarial_cloud(n, seed=0, dtype=np.float32):
    """Tie-heavy input: grid-snapped coordinates, 10 % exact duplicates and a cluster
    inside the FPS skip ball (|p|^2 <= 1e-3)."""
    rng = np.random.default_rng(seed)
    pts = np.round(rng.uniform(-0.3, 0.3, (n, 3)) * 50) / 50
    pts[:, 2] += 0.6
    dup = rng.choice(n, n // 10, replace=False)
    pts[dup] = pts[rng.choice(n, n // 10)]
    near0 = rng.choice(n, max(n // 50, 1), replace=False)
    pts[near0] = rng.uniform(-0.015, 0.015, (len(near0), 3))
    return np.ascontiguousarray(pts, dtype=dtype)


def fibonacci_views(num_views=300):
    """Unit view directions on a Fibonacci sphere lattice."""
    i = np.arange(num_views)
    phi = (np.sqrt(5) - 1) / 2
    zi = (2 * i + 1) / num_views - 1
    r = np.sqrt(np.clip(1 - zi ** 2, 0, None))
    return np.stack([r * np.cos(2 * i * np.pi * phi), r * np.sin(2 * i * np.pi * phi), zi], 1)


def view_rotations(towards, angle):
    """Proper rotations whose first column is `towards` (unit, (K,3)), rotated in-plane by
    `angle` (K,) about that axis.  Row-major (K,3,3) fp64."""
    ax = towards / np.linalg.norm(towards, axis=1, keepdims=True)
    ay = np.stack([-ax[:, 1], ax[:, 0], np.zeros(len(ax))], 1)
    bad = np.linalg.norm(ay, axis=1) < 1e-9
    ay[bad] = np.array([0.0, 1.0, 0.0])
    ay /= np.linalg.norm(ay, axis=1, keepdims=True)
    az = np.cross(ax, ay)
    base = np.stack([ax, ay, az], 2)
    c, s = np.cos(angle), np.sin(angle)
    z, o = np.zeros_like(c), np.ones_like(c)
    r1 = np.stack([np.stack([o, z, z], 1), np.stack([z, c, -s], 1), np.stack([z, s, c], 1)], 1)
    return base @ r1


def seed_rotations(k, seed=0, dtype=np.float32):
    """(k,3,3) approach-frame rotations as ApproachNet would emit (view from the 300-view
    lattice, pointing roughly towards +z = away from the camera, angle 0)."""
    rng = np.random.default_rng(seed)
    v = fibonacci_views(300)
    v = v[v[:, 2] > 0.2]
    t = v[rng.integers(0, len(v), k)]
    return np.ascontiguousarray(view_rotations(t, np.zeros(k)), dtype=dtype)


def synthetic_grasps(scene, m, seed=0, dtype=np.float32):
    """(m,17) rows [score,width,height,depth,R(9),t(3),obj]: score U(0,1), width U(0,0.1),
    height 0.02, depth in {0.01..0.04}, R from lattice view + angle k*pi/12, t = scene points."""
    rng = np.random.default_rng(seed)
    g = np.zeros((m, 17), np.float64)
    g[:, 0] = rng.uniform(0, 1, m)
    g[:, 1] = rng.uniform(0, 0.1, m)
    g[:, 2] = 0.02
    g[:, 3] = rng.integers(1, 5, m) * 0.01
    v = fibonacci_views(300)
    v = v[v[:, 2] > 0.2]
    t = v[rng.integers(0, len(v), m)]
    g[:, 4:13] = view_rotations(t, rng.integers(0, 12, m) * np.pi / 12).reshape(m, 9)
    g[:, 13:16] = np.asarray(scene, dtype=np.float64)[rng.choice(len(scene), m, replace=len(scene) < m)]
    g[:, 16] = -1
    return np.ascontiguousarray(g, dtype=dtype)


def depth_to_cloud(depth, intrinsic, factor_depth):
    """(H,W) depth image -> (H,W,3) float64 camera-frame points: z = depth / factor, x = (u - cx) z / fx,
    y = (v - cy) z / fy (the pinhole back-projection of utils/data_utils.py:18-43)"""
    h, w = depth.shape
    u, v = np.meshgrid(np.arange(w), np.arange(h))
    z = depth / factor_depth
    x = (u - intrinsic[0, 2]) * z / intrinsic[0, 0]
    y = (v - intrinsic[1, 2]) * z / intrinsic[1, 1]
    return np.stack([x, y, z], axis=-1)


def fixture_scene(path):
    """The reference's example scene (doc/example_data: depth.png, meta.mat, workspace_mask.png) from the committed
    fixture tests/golden/scene_example.npz -> (points (513688,3) float64 in image order, fixture dict).  The file holds
    the depth image cropped to the workspace mask, the mask and the intrinsics (tests/golden/make_golden.py scene)."""
    g = np.load(path)
    y0, y1, x0, x1 = [int(v) for v in g["crop"]]
    h, w = [int(v) for v in g["shape"]]
    depth = np.zeros((h, w), np.uint16)
    depth[y0:y1, x0:x1] = g["depth_crop"]
    mask = np.zeros((h, w), bool)
    mask[y0:y1, x0:x1] = np.unpackbits(g["mask_crop"])[: (y1 - y0) * (x1 - x0)].reshape(y1 - y0, x1 - x0).astype(bool)
    cloud = depth_to_cloud(depth, g["intrinsic"], float(g["factor_depth"]))
    return cloud[mask], g


def training_labels(clouds, num_objects=5, num_grasp_points=1200, num_view=300, num_angle=12, num_depth=4, seed=0,
                    device="cpu"):
    """Synthetic labels of a training batch with the keys and shapes graspnet_dataset.py:238-246 produces (SURVEY.md
    8(d) item 5): clouds (B,N,3) numpy -> dict of 'objectness_label' (B,N) i64 and per-scene lists (one entry per
    object) of 'object_poses_list' (3,4), 'grasp_points_list' (Np,3), 'grasp_offsets_list' (Np,V,A,D,3) [angle, depth,
    width], 'grasp_labels_list' (Np,V,A,D) friction coefficients U(0,1.2) with 70 % zeros (and a few -1),
    'grasp_tolerance_list' (Np,V,A,D).  Objects sit on points of the cloud so that seeds find near grasp points."""
    import torch
    rng = np.random.default_rng(seed)
    B, N, _ = clouds.shape
    out = {k: [] for k in ("object_poses_list", "grasp_points_list", "grasp_offsets_list", "grasp_labels_list",
                           "grasp_tolerance_list")}
    obj_label = np.zeros((B, N), np.int64)
    f = np.float32
    V, A, D = num_view, num_angle, num_depth
    for b in range(B):
        scene = {k: [] for k in out}
        centres = clouds[b][rng.choice(N, num_objects, replace=False)]
        near = np.linalg.norm(clouds[b][:, None, :] - centres[None], axis=2).min(axis=1) < 0.08
        obj_label[b] = near
        for o in range(num_objects):
            q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
            if np.linalg.det(q) < 0:
                q[:, 0] = -q[:, 0]
            pose = np.concatenate([q, centres[o][:, None]], axis=1).astype(f)
            pts = (rng.standard_normal((num_grasp_points, 3)) * 0.03).astype(f)           # object frame
            lab = rng.uniform(0.0, 1.2, (num_grasp_points, V, A, D)).astype(f)
            r = rng.random(lab.shape)
            lab[r < 0.7] = 0.0
            lab[r > 0.995] = -1.0
            off = np.empty((num_grasp_points, V, A, D, 3), f)
            off[..., 0] = (np.arange(A, dtype=f) * (np.pi / A))[None, None, :, None]
            off[..., 1] = (np.arange(1, D + 1, dtype=f) * 0.01)[None, None, None, :]
            off[..., 2] = rng.uniform(0.0, 0.12, lab.shape).astype(f)
            tol = rng.uniform(0.0, 0.05, lab.shape).astype(f)
            for k, v in (("object_poses_list", pose), ("grasp_points_list", pts), ("grasp_offsets_list", off),
                         ("grasp_labels_list", lab), ("grasp_tolerance_list", tol)):
                scene[k].append(torch.from_numpy(v).to(device))
        for k in out:
            out[k].append(scene[k])
    out["objectness_label"] = torch.from_numpy(obj_label).to(device)
    return out


def training_predictions(seed, b, ns, v=300, a=12, d=4):
    """seeded stand-ins for the network outputs the loss reads (numpy; the tests rebuild the same values)"""
    rng = np.random.default_rng(seed)
    f = np.float32
    return {"objectness_score": rng.standard_normal((b, 2, ns)).astype(f),
            "view_score": rng.uniform(0, 1.5, (b, ns, v)).astype(f),
            "grasp_top_view_inds": rng.integers(0, v, (b, ns)).astype(np.int64),
            "grasp_top_view_rot": np.zeros((b, ns, 3, 3), f),       # read but unused by models/loss.py:71
            "grasp_score_pred": rng.uniform(0, 2, (b, a, ns, d)).astype(f),
            "grasp_angle_cls_pred": rng.standard_normal((b, a, ns, d)).astype(f),
            "grasp_width_pred": rng.uniform(0, 0.1, (b, a, ns, d)).astype(f),
            "grasp_tolerance_pred": rng.uniform(0, 0.05, (b, a, ns, d)).astype(f)}


def mlp_params(seed, dims):
    """Seeded parameters of a shared MLP (SharedMLP(dims, bn=True), reference pytorch_utils.py:11-36):
    [(conv weight (Cout,Cin), bn gamma, beta, running mean, running var)] per layer.  numpy only, so the golden
    generator and the tests build exactly the same values."""
    rng = np.random.default_rng(seed)
    out = []
    for cin, cout in zip(dims[:-1], dims[1:]):
        w = (rng.standard_normal((cout, cin)) * np.sqrt(2.0 / cin)).astype(np.float32)
        out.append((w, rng.uniform(0.5, 1.5, cout).astype(np.float32),
                    (0.2 * rng.standard_normal(cout)).astype(np.float32),
                    (0.3 * rng.standard_normal(cout)).astype(np.float32),
                    rng.uniform(0.5, 2.0, cout).astype(np.float32)))
    return out


def fill_state_dict(module, seed):
    """Deterministic parameters and BatchNorm statistics for a torch module, keyed by TENSOR NAME (so two
    implementations with the same state_dict keys -- the reference GraspNet and this package's -- get the same
    values whatever their construction order): conv weights N(0, 2/fan_in), biases and running means N(0, 0.1),
    BatchNorm scales and running variances U(0.5, 1.5)."""
    import zlib
    import torch
    sd = module.state_dict()
    with torch.no_grad():
        for name in sorted(sd):
            t = sd[name]
            if name.endswith("num_batches_tracked"):
                continue
            rng = np.random.default_rng([seed, zlib.crc32(name.encode())])
            shape = tuple(t.shape)
            if name.endswith("running_var") or (t.dim() == 1 and name.endswith("weight")):
                v = rng.uniform(0.5, 1.5, shape)
            elif name.endswith("running_mean") or name.endswith("bias"):
                v = rng.normal(0.0, 0.1, shape)
            else:
                v = rng.normal(0.0, np.sqrt(2.0 / max(int(np.prod(shape[1:])), 1)), shape)
            t.copy_(torch.from_numpy(v.astype(np.float32)).to(t.device))
    return module
