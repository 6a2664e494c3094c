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
