"""The inequality behind the drop gates of the primary kernel (kernels.cu, sky_drop_distance), held
on the CPU against the plain fp32 DDA it makes a promise about.  A ray that steps DOWN may enter free flight when its y, less
an upper bound on what it can still lose, clears the column tops it can still reach.  The bound has two parts:
  total drop : of the `left` events still to fire at most floor((left + 3) * |dy| / (|dx| + |dy| + |dz|)) + 2 are y events;
  ring drop  : while fewer than D = (d + 1) * 16 steps have been taken along x and along z (the ray is still within d
               column blocks of its own), at most floor(D * min(absDelta.x, absDelta.z) * |dy|) + 2 y events have fired.
This test restates the kernel's fp32 arithmetic for both and walks nextIntersectDDA (src/shaders/low_res.frag:446-491:
the smallest deltaPos fires, x if tx < ty && tx < tz, else y if ty < tz, else z; one rounded fp32 addition) for hundreds of
thousands of rays -- random, grazing, axis-parallel, diagonal ties, start fractions on voxel faces -- from random moments of
their flight (the gates are asked in mid-flight), and checks that no ray ever loses more height than promised."""
import numpy as np

F = np.float32


def _rays(rng, n):
    d = rng.normal(size=(n, 3))
    kind = rng.integers(0, 8, size=n)
    g = kind == 0                      # grazing: the rays the gates exist for
    d[g, 1] = -np.abs(rng.uniform(1e-4, 0.08, size=g.sum())) * np.linalg.norm(d[g][:, [0, 2]], axis=1)
    a = kind == 1                      # one horizontal axis never steps (absDelta = inf)
    d[a, 0] = 0.0
    b = kind == 2                      # the other one
    d[b, 2] = 0.0
    t = kind == 3                      # exact diagonals: ties on every step
    d[t, 0] = np.sign(d[t, 0]) * np.abs(d[t, 2])
    s = kind == 4                      # steep
    d[s, 1] *= 20.0
    d[:, 1] = -np.abs(d[:, 1]) - 1e-9  # all step down
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(F)
    frac = rng.uniform(0, 1, size=(n, 3)).astype(F)
    snap = rng.random(size=(n, 3)) < 0.15
    frac[snap] = rng.choice(np.array([0.0, 0.5, 1.0 - 2.0 ** -24], dtype=F), size=int(snap.sum()))
    return d, frac


def _build(d, frac):
    """buildRay + the initial deltaPos of low_res.frag:61-97,296 in fp32 (absDelta = |1 / dir|, deltaPos = absDelta -
    (exact - round) * delta with exact biased by -1 for negative steps)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        delta = (F(1.0) / d).astype(F)
        ad = np.abs(delta)
        ex = np.where(d < 0, (frac - F(1.0)).astype(F), frac)
        t = (ad - (ex * delta).astype(F)).astype(F)
    return ad, t


def test_drop_bounds_hold_for_the_fp32_dda():
    rng = np.random.default_rng(20261105)
    n, steps = 150_000, 1024
    d, frac = _rays(rng, n)
    ad, t = _build(d, frac)
    ok = ~np.isnan(t).any(axis=1)      # the kernel asks the gates only when no deltaPos is NaN
    d, ad, t = d[ok], ad[ok], t[ok]
    n = len(d)
    # the moment the gate is asked: after `start` steps; budget `left` from there
    start = rng.integers(0, 200, size=n)
    left = rng.choice(np.array([40, 100, 300, 500, 1024]), size=n)
    left = np.minimum(left, steps - start)
    with np.errstate(over="ignore", invalid="ignore"):
        rate = (np.abs(d[:, 1]) / ((np.abs(d[:, 0]) + np.abs(d[:, 1])).astype(F) + np.abs(d[:, 2])).astype(F)).astype(F)
        total = np.floor(((left + 3).astype(F) * rate).astype(F)) + F(2.0)
        slope = (np.minimum(ad[:, 0], ad[:, 2]) * np.abs(d[:, 1])).astype(F) * F(16.0)
    rings = {}
    for k in (2, 4, 8, 16):            # (d + 1) of rings 0-1, 2-3, 4-7, 8-15
        with np.errstate(invalid="ignore", over="ignore"):
            r = np.floor(slope * F(k)) + F(2.0)
        rings[k] = np.where(np.isnan(r), total, np.minimum(total, r))   # fminf(total, NaN) = total
    cnt = np.zeros((n, 3), dtype=np.int64)       # events since the gate was asked
    worst_total = np.zeros(n, dtype=np.int64)
    worst_ring = {k: np.zeros(n, dtype=np.int64) for k in rings}
    idx = np.arange(n)
    for step in range(steps):
        tx, ty, tz = t[:, 0], t[:, 1], t[:, 2]
        ax = np.where((tx < ty) & (tx < tz), 0, np.where(ty < tz, 1, 2))
        t[idx, ax] = (t[idx, ax] + ad[idx, ax]).astype(F)
        live = (step >= start) & (step < start + left)
        cnt[idx[live], ax[live]] += 1
        worst_total = np.where(live, cnt[:, 1], worst_total)
        for k in rings:
            inside = live & (cnt[:, 0] <= 16 * k - 1) & (cnt[:, 2] <= 16 * k - 1)
            worst_ring[k] = np.where(inside, np.maximum(worst_ring[k], cnt[:, 1]), worst_ring[k])
    assert (worst_total <= total).all(), np.flatnonzero(worst_total > total)[:5]
    for k in rings:
        bad = worst_ring[k] > rings[k]
        assert not bad.any(), (k, np.flatnonzero(bad)[:5], worst_ring[k][bad][:5], rings[k][bad][:5], d[bad][:5])
    # the bounds are not vacuous: grazing rays are promised (and lose) only a few voxels over 300 steps
    grazing = (np.abs(d[:, 1]) < 0.03) & (left == 300)
    assert grazing.sum() > 1000 and np.median(total[grazing]) <= 8 and (worst_total[grazing] <= 8).mean() > 0.9
    # ... and tight enough to matter: the slack over what really happened is 2-3 events for most rays
    assert np.median(total - worst_total) <= 3


def test_climb_credit_of_the_sky_gates_holds_for_the_fp32_dda():
    """The mirror image, for rays that step UP (sky_gate_distance): a ring whose nearest column block is d blocks away
    (d = 2, 4, 8, 16) is credited with the height the ray must have gained before it can be there,
    max(floor(min(absDelta.x, absDelta.z) * |dy| * 16 * (d - 1)) - 3, 0) voxels (the per-voxel slope clamped to [0, 16]).
    To be d blocks away the ray has taken at least (d - 1) * 16 + 1 steps along x or along z."""
    rng = np.random.default_rng(77)
    n, steps = 150_000, 1024
    d, frac = _rays(rng, n)
    d[:, 1] = -d[:, 1]                 # all step up
    ad, t = _build(d, frac)
    ok = ~np.isnan(t[:, 1]) & ~np.isnan(t[:, 2])   # merge = false (NaN on y or z) gets no credit in the kernel
    d, ad, t = d[ok], ad[ok], t[ok]
    n = len(d)
    start = rng.integers(0, 200, size=n)
    with np.errstate(invalid="ignore", over="ignore"):
        per_voxel = np.minimum(np.maximum((np.minimum(ad[:, 0], ad[:, 2]) * np.abs(d[:, 1])).astype(F), F(0.0)), F(16.0))
        per_voxel = np.where(np.isnan(per_voxel), F(0.0), per_voxel)   # fmaxf / fminf turn a NaN product into 0
        slope = (per_voxel * F(16.0)).astype(F)
    credit = {k: np.maximum(np.floor((slope * F(k)).astype(F)) - 3, 0) for k in (1, 3, 7, 15)}
    cnt = np.zeros((n, 3), dtype=np.int64)
    reached = {k: np.full(n, -1, dtype=np.int64) for k in credit}   # y events when the ring is first reachable
    idx = np.arange(n)
    for step in range(steps):
        tx, ty, tz = t[:, 0], t[:, 1], t[:, 2]
        ax = np.where((tx < ty) & (tx < tz), 0, np.where(ty < tz, 1, 2))
        t[idx, ax] = (t[idx, ax] + ad[idx, ax]).astype(F)
        live = step >= start
        cnt[idx[live], ax[live]] += 1
        far = np.maximum(cnt[:, 0], cnt[:, 2])
        for k in credit:
            first = live & (reached[k] < 0) & (far >= 16 * k + 1)
            reached[k] = np.where(first, cnt[:, 1], reached[k])
    seen = 0
    for k in credit:
        m = reached[k] >= 0
        seen += int(m.sum())
        bad = m & (reached[k] < credit[k])
        assert not bad.any(), (k, np.flatnonzero(bad)[:5], reached[k][bad][:5], credit[k][bad][:5], d[bad][:5])
    assert seen > n  # most rays get to the near rings within the steps walked
    # not vacuous: steep rays are credited with real height
    assert (credit[15] > 100).mean() > 0.1
