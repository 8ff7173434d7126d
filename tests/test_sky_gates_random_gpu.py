"""Randomised hold on the two mechanisms that let a primary ray stop asking the tree: free flight (sky gates + the closed-form
leap, kernels.cu loop A, csrc/free_flight.h) and the column regions (everything above a column block's top is one empty
region).  Both rest on the per-column-block tops of the world, and on inequalities (ring distances, slope credit with its
"- 3" slack) that hand-picked cameras only sample.  On-versus-off needs no oracle and costs microseconds per frame, so:
2,400 random cases = 120 random sparse worlds (columns, floating single voxels / 4^3 / 16^3 cells, thin spikes, cells that
straddle the borders of the 16 x 16 column blocks and the torus seam; tree depths 6 and 7) x 20 random (camera, pitch up and
down, step budget 40..1024, sun incl. near-horizontal and below the horizon) -- free flight on / walked / off, column
regions on / off: hit records (final voxel, face and step count of EVERY ray, sky rays included) and RGBA8 identical; one
case per world also against the oracle.  A failure prints the seed that reproduces it."""
import numpy as np
import pytest

from conftest import camera_dir

pytestmark = pytest.mark.gpu

N_WORLDS, N_CASES = 120, 20
BUDGETS = [40, 64, 100, 300, 300, 300, 500, 1000, 1024]


def random_world(vh, rng, depth):
    host = vh.World(depth)
    edge = 1 << (2 * (depth - 1))
    block = 16 if depth == 6 else 64          # column-block edge of the sky grid (64 x 64 blocks over the torus)
    base = 0 if depth == 6 else int(rng.integers(0, 3)) * 1024
    span = int(rng.choice([64, 160, 400]))
    col = lambda: int(rng.integers(1, 1 << 40))

    def border(v):  # snap some coordinates onto / next to a column-block border
        v = int(v // block * block + rng.choice([-1, 0, 1])) if rng.random() < 0.3 else int(v)
        return min(max(v, 0), edge - 1)

    if rng.random() < 0.6:                     # a ground to start shadow / AO rays from
        gy = int(rng.integers(0, 6))
        step = int(rng.choice([2, 3, 4]))
        for x in range(0, span, step):
            for z in range(0, span, step):
                host.put_block(base + x, gy, base + z, col(), 0, 0.0, depth - 1 if step == 4 else depth)
    for _ in range(int(rng.integers(5, 60))):  # columns of random height
        x, z = border(base + rng.integers(0, span)), border(base + rng.integers(0, span))
        top = int(rng.integers(1, 150))
        for y in range(max(0, top - int(rng.integers(1, 6))), top + 1):
            host.put_block(x, y, z, col(), 0, 0.0)
    for _ in range(int(rng.integers(0, 25))):  # floating single voxels, some mirrors / refractors
        host.put_block(border(base + rng.integers(0, span)), int(rng.integers(5, 300)), border(base + rng.integers(0, span)), col(),
                       int(rng.choice([0, 0, 0, 2, 4, 0x14])), 0.0)
    for _ in range(int(rng.integers(0, 12))):  # floating 4^3 cells, aligned, often next to a block border
        x, z = border(base + rng.integers(0, span)) // 4 * 4, border(base + rng.integers(0, span)) // 4 * 4
        host.put_block(x, int(rng.integers(1, 60)) * 4, z, col(), 0, 0.0, depth - 1)
    for _ in range(int(rng.integers(0, 4))):   # 16^3 cells
        host.put_block(base + int(rng.integers(0, max(span // 16, 1))) * 16, int(rng.integers(0, 12)) * 16,
                       base + int(rng.integers(0, max(span // 16, 1))) * 16, col(), 0, 0.0, depth - 2)
    if rng.random() < 0.4:                     # something at the torus seam: the rings run off the grid there
        for _ in range(int(rng.integers(1, 6))):
            x = int(rng.choice([0, 1, edge - 2, edge - 1]))
            host.put_block(x, int(rng.integers(0, 120)), int(rng.choice([0, 3, edge - 1, base + 20])), col(), 0, 0.0)
    return host, base, span, edge


def random_case(rng, base, span, edge):
    kind = rng.random()
    if kind < 0.7:    # inside / above the populated part
        o = (base + rng.uniform(-10, span + 10), rng.uniform(0.2, 180.0), base + rng.uniform(-10, span + 10))
    elif kind < 0.85:  # near the seam (negative coordinates at depth 6 go through the reference's Morton quirk)
        o = (rng.choice([-20.5, -0.5, 0.5, edge - 1.5, edge + 3.5]) + 0.0, rng.uniform(0.2, 120.0), base + rng.uniform(0, span))
    else:             # high up: y + budget may leave the torus along y
        o = (base + rng.uniform(0, span), rng.uniform(200.0, edge - 1.0), base + rng.uniform(0, span))
    o = tuple(float(np.float32(v)) for v in o)
    if rng.random() < 0.1:  # axis-parallel: absDelta = inf, NaN deltaPos on a voxel corner
        d = [(1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, 0.0, -1.0), (0.0, -1.0, 0.0)][int(rng.integers(0, 4))]
        if rng.random() < 0.5:
            o = (float(np.floor(o[0])), float(np.floor(o[1])), float(np.floor(o[2])))
    else:
        d = camera_dir(float(rng.uniform(-180, 180)), float(rng.choice([rng.uniform(-89, 89), rng.uniform(-12, 3)])))
    s = np.array([rng.normal(), rng.choice([rng.uniform(0.01, 0.2), rng.uniform(0.2, 2.0), -rng.uniform(0.05, 1.0)], p=[0.35, 0.5, 0.15]),
                  rng.normal()], dtype=np.float32)
    s = s / np.float32(np.sqrt(np.float32((s * s).sum())))
    return o, d, tuple(float(v) for v in s), int(rng.choice(BUDGETS))


def test_sky_gates_random_worlds(ob, rcmod, caster, world_desc):
    from raytracing_test_b200 import voxel_host as vh

    modes = {"plain": dict(free_flight=False, column_regions=False), "on": {}, "walk": dict(leap=False),
             "no_columns": dict(column_regions=False), "no_free_flight": dict(free_flight=False), "no_drop_gates": dict(drop_gates=False)}
    W, H = 160, 90
    free = cases = oracle_cases = 0
    try:
        for wi in range(N_WORLDS):
            seed = 7_000_000 + wi
            rng = np.random.default_rng(seed)
            depth = 6 if wi % 3 else 7
            host, base, span, edge = random_world(vh, rng, depth)
            try:
                caster.upload_world(host.describe())
                oworld = None
                for ci in range(N_CASES):
                    o, d, sun, budget = random_case(rng, base, span, edge)
                    shadow_steps = int(rng.choice([75, min(budget, 75), budget]))
                    kw = dict(shadows=True, max_steps=budget, shadow_steps=shadow_steps, hits=True)
                    got = {}
                    for name, m in modes.items():
                        caster.render_frame(rcmod.frame_params(W, H, o, d, sun, **kw, **m))
                        got[name] = (caster.read_image().copy(), caster.read_hits().copy())
                        if name == "on":
                            free += caster.stats().free_flight_rays
                    where = f"seed {seed} (world {wi}, depth {depth}) case {ci}: origin {o} dir {tuple(float(v) for v in d)} sun {sun} " \
                            f"max_steps {budget} shadow_steps {shadow_steps}"
                    for name in modes:
                        if name == "plain":
                            continue
                        a, b = got[name], got["plain"]
                        if a[1].tobytes() != b[1].tobytes():
                            bad = np.flatnonzero((a[1].view(np.uint8).reshape(-1, 20) != b[1].view(np.uint8).reshape(-1, 20)).any(1))
                            raise AssertionError(f"{where}: mode {name}: {len(bad)} hit records differ, first pixel {bad[0]}: "
                                                 f"{a[1][bad[0]]} vs plain {b[1][bad[0]]}")
                        assert np.array_equal(a[0], b[0]), f"{where}: mode {name}: RGBA8 differs"
                    cases += 1
                    if ci == 0:  # one case per world against the oracle too (the tree-only walk is then pinned as well)
                        if oworld is None:
                            nodes, arrays = host.pools()
                            oworld = ob.make_world(nodes, arrays, depth)
                        ref = ob.render(oworld, ob.make_frame(W, H, o, d, sun, shadows=True, max_steps=budget, shadow_steps=shadow_steps))
                        assert got["on"][1].tobytes() == ref["hits"].tobytes(), f"{where}: differs from the oracle"
                        assert np.array_equal(got["on"][0], ref["rgba8"]), f"{where}: RGBA8 differs from the oracle"
                        oracle_cases += 1
            finally:
                host.close()
    finally:
        caster.upload_world(world_desc)
    assert cases == N_WORLDS * N_CASES >= 2000 and oracle_cases == N_WORLDS
    assert free > 100 * cases  # free flight really runs in these worlds (most rays that look up)
