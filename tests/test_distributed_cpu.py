"""Host-side logic of the multi-GPU paths (ovr_b200.distributed), on CPU: brick decomposition, visibility order, the
binary-swap schedule run for real over torch.distributed/gloo with world_size 2 and 4, and the sort-first slice layout.
No rendering happens here: the per-rank images are synthetic premultiplied layers and the blend is the injected
`over_torch`, so what is tested is who sends which half to whom and who is composited in front."""
import os
import socket

import numpy as np
import pytest

import ovr_b200
from ovr_b200 import distributed as D


def test_brick_grid_and_indices_partition_the_volume():
    for n in (1, 2, 4, 8, 16):
        g = D.brick_grid(n)
        assert g[0] * g[1] * g[2] == n
        idx = {D.brick_index(r, n) for r in range(n)}
        assert idx == {(i, j, k) for i in range(g[0]) for j in range(g[1]) for k in range(g[2])}
    with pytest.raises(ValueError):
        D.brick_grid(6)
    dims = (50, 64, 33)   # odd sizes: cuts at floor(dims * i / g)
    owned = np.zeros(dims, np.int32)
    for r in range(8):
        b = D.Brick(dims, r, 8)
        owned[b.own_lo[0]:b.own_hi[0], b.own_lo[1]:b.own_hi[1], b.own_lo[2]:b.own_hi[2]] += 1
        for a in range(3):
            # ghosts: one voxel below, two above every interior cut, none outside the volume
            assert b.org[a] == max(b.own_lo[a] - 1, 0) and b.org[a] + b.size[a] == min(b.own_hi[a] + 2, dims[a])
            # boxes in brick object space map back to the right voxels
            assert np.isclose(float(b.clip_lo[a]) * b.size[a] + b.org[a], b.own_lo[a], atol=1e-4)
            assert np.isclose(float(b.clip_hi[a]) * b.size[a] + b.org[a], b.own_hi[a], atol=1e-4)
            assert b.own_lo_w[a] == b.own_lo[a] and b.own_hi_w[a] == b.own_hi[a]   # spacing 1, origin 0
            assert b.vol_lo_w[a] == 0 and b.vol_hi_w[a] == dims[a]
        # unit cube of the brick lands on its stored voxels in world space (spacing 1, origin 0)
        assert np.allclose(b.obj_to_world[:, 3], b.org) and np.allclose(np.diag(b.obj_to_world[:, :3]), b.size)
    assert (owned == 1).all()


def test_brick_scene_slices_data_and_offsets_noise():
    sc = ovr_b200.scenes.noise_scene(32, tf="ramp", camera="oblique")
    if sc.tfn.value_range[0] > sc.tfn.value_range[1]:
        sc.tfn.value_range = (0.0, 1.0)
    part, b = D.brick_scene(sc, 3, 4)
    assert part.volume.noise["dims"] == b.size and tuple(part.volume.noise["org"]) == b.org
    assert np.allclose(part.volume.obj_to_world().reshape(3, 4), b.obj_to_world)
    data = np.arange(32 * 32 * 32, dtype=np.float32).reshape(32, 32, 32)
    sc2 = ovr_b200.Scene(ovr_b200.Volume(data=data), sc.tfn, sc.camera, 1.0)
    part2, b2 = D.brick_scene(sc2, 3, 4)
    (x0, y0, z0), (sx, sy, sz) = b2.org, b2.size
    assert np.array_equal(part2.volume.data, data[z0:z0 + sz, y0:y0 + sy, x0:x0 + sx])
    bad = ovr_b200.Scene(sc.volume, ovr_b200.TransferFunction(sc.tfn.color, sc.tfn.opacity, (1.0, -1.0)), sc.camera, 1.0)
    with pytest.raises(ValueError):
        D.brick_scene(bad, 0, 2)


def test_visibility_order_matches_a_brute_force_ray_walk():
    """For random eyes, the per-round front/back decision must agree with the order in which a ray from the eye meets
    the two boxes being merged in that round."""
    rng = np.random.default_rng(3)
    dims = (64, 64, 64)
    for n in (2, 4, 8, 16):
        g = D.brick_grid(n)
        for _ in range(20):
            eye = rng.uniform(-100, 164, 3)
            for rank in range(n):
                for bit in range(n.bit_length() - 1):
                    axis, plane = D.split_plane(dims, n, bit, rank)
                    assert axis == bit % 3
                    # at round `bit` a rank holds the composite of the bricks that agree with it in all bits >= bit
                    def group_box(r):
                        members = [m for m in range(n) if (m >> bit) == (r >> bit)]
                        lo = np.min([D.Brick(dims, m, n).own_lo for m in members], axis=0)
                        hi = np.max([D.Brick(dims, m, n).own_hi for m in members], axis=0)
                        return lo, hi
                    lo_a, hi_a = group_box(rank)
                    lo_b, hi_b = group_box(rank ^ (1 << bit))
                    # the two group boxes touch exactly at the split plane and agree on the other two axes
                    assert (hi_a[axis] == plane and lo_b[axis] == plane) or (hi_b[axis] == plane and lo_a[axis] == plane)
                    for a in range(3):
                        if a != axis:
                            assert lo_a[a] == lo_b[a] and hi_a[a] == hi_b[a]
                    if eye[axis] == plane:
                        continue
                    # the group on the eye's side of the plane is met first by every ray from the eye
                    want_front = (eye[axis] < plane) == (0.5 * (lo_a[axis] + hi_a[axis]) < plane)
                    assert D.in_front(eye, dims, n, bit, rank) == want_front
                    assert D.in_front(eye, dims, n, bit, rank) != D.in_front(eye, dims, n, bit, rank ^ (1 << bit))


def test_binary_swap_plan_ranges_tile_the_image():
    for n in (1, 2, 4, 8):
        for npix in (1920 * 1080, 1000, 37 * 32):
            ranges = D.final_ranges(n, npix, align=32)
            assert sorted(ranges)[0][0] == 0 and sorted(ranges)[-1][1] == npix
            covered = np.zeros(npix, np.int32)
            for lo, hi in ranges:
                covered[lo:hi] += 1
            assert (covered == 1).all()
            for r in range(n):
                plan, _ = D.binary_swap_plan(r, n, npix, 32)
                for b, (partner, keep, send) in enumerate(plan):
                    pplan, _ = D.binary_swap_plan(partner, n, npix, 32)
                    assert pplan[b][0] == r and pplan[b][1] == send and pplan[b][2] == keep   # halves are exchanged


def test_assemble_slices_matches_the_round_robin_layout():
    W, H, n = 200, 120, 3
    frame = np.random.default_rng(0).random((H, W, 4), dtype=np.float32)
    sw, sh = D.super_tile_grid(W, H)
    slices = []
    for rank in range(n):
        tiles = D.owned_super_tiles(W, H, rank, n)
        sl = np.zeros((len(tiles), 32, 64, 4), np.float32)
        for m, t in enumerate(tiles):
            x0, y0 = (t % sw) * 64, (t // sw) * 32
            w, h = min(64, W - x0), min(32, H - y0)
            sl[m, :h, :w] = frame[y0:y0 + h, x0:x0 + w]
        assert sl.size == D.tile_slice_floats(W, H, rank, n)
        slices.append(sl.ravel())
    assert np.array_equal(D.assemble_slices(slices, W, H), frame)
    assert sorted(sum((D.owned_super_tiles(W, H, r, n) for r in range(n)), [])) == list(range(sw * sh))


# ---- the schedule over a real process group ------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _layer(rank, npix, seed=11):
    """Synthetic premultiplied layer of rank's brick: alpha in [0, 0.9], colour <= alpha."""
    rng = np.random.default_rng(seed + rank)
    a = rng.uniform(0.0, 0.9, (npix, 1)).astype(np.float32)
    c = (rng.uniform(0.0, 1.0, (npix, 3)).astype(np.float32)) * a
    return np.concatenate([c, a], axis=1)


def _worker(rank, world, port, npix, eye, dims, out_q, schedule="swap"):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cls = D.SortLast if schedule == "swap" else D.SortLastDirect
        sl = cls(rank, world, dims, npix, blend=D.over_torch, align=8)
        img = torch.from_numpy(_layer(rank, npix).ravel().copy())
        with pytest.raises(RuntimeError):   # no injected blend + CPU tensors: the product path refuses
            D.SortLast(rank, world, dims, npix)._over(img[:8], img[:8], img[:8].clone())
        res = sl.composite(img, eye)
        if rank == 0:
            out_q.put(res.numpy().reshape(npix, 4).copy())
    finally:
        dist.destroy_process_group()


def _expected(world, npix, eye, dims):
    """Composite the layers in true visibility order: sort bricks front to back along every axis by the eye's side."""
    layers = {r: _layer(r, npix) for r in range(world)}
    g = D.brick_grid(world)

    def key(r):   # eye outside the volume: along every axis the bricks nearer to the eye come first
        idx = D.brick_index(r, world)
        return tuple(idx[a] if eye[a] < 0 else g[a] - 1 - idx[a] for a in (2, 1, 0))
    out = np.zeros((npix, 4), np.float32)
    for r in sorted(range(world), key=key):
        out = out + (1.0 - out[:, 3:4]) * layers[r]
    return out


def test_direct_send_order_and_ranges():
    """visibility_order: a total order that agrees with the brute-force per-axis sort for eyes outside the volume, and with
    every round of the binary swap for any eye; direct_send_ranges tile the image."""
    rng = np.random.default_rng(9)
    dims = (64, 48, 80)
    for n in (1, 2, 4, 8, 16):
        g = D.brick_grid(n)
        for _ in range(25):
            eye = rng.uniform(-150, 230, 3)
            order = D.visibility_order(eye, dims, n)
            assert sorted(order) == list(range(n))
            pos = {r: i for i, r in enumerate(order)}
            for r in range(n):
                for bit in range(n.bit_length() - 1):
                    # at round `bit` of the swap r's group is blended in front of its partner's group iff in_front says so
                    p = r ^ (1 << bit)
                    mine = [m for m in range(n) if (m >> bit) == (r >> bit)]
                    theirs = [m for m in range(n) if (m >> bit) == (p >> bit)]
                    if D.in_front(eye, dims, n, bit, r):
                        assert max(pos[m] for m in mine) < min(pos[m] for m in theirs)
            outside = all(eye[a] < 0 or eye[a] > dims[a] for a in range(3))
            if outside:   # along every axis nearer bricks first: any order consistent with that partial order is valid
                for a_ in range(n):
                    for b_ in range(n):
                        ia, ib = D.brick_index(a_, n), D.brick_index(b_, n)
                        nearer = all((ia[x] <= ib[x]) if eye[x] < 0 else (ia[x] >= ib[x]) for x in range(3))
                        if nearer and a_ != b_:
                            assert pos[a_] < pos[b_]
        for npix in (1920 * 1080, 1000, 37 * 32, 5):
            ranges = D.direct_send_ranges(n, npix, align=32)
            covered = np.zeros(npix, np.int32)
            for lo, hi in ranges:
                assert 0 <= lo <= hi <= npix and (lo % 32 == 0 or lo == npix)
                covered[lo:hi] += 1
            assert (covered == 1).all() and len(ranges) == n


@pytest.mark.parametrize("world,schedule", [(2, "swap"), (4, "swap"), (2, "direct"), (4, "direct")])
def test_binary_swap_over_gloo_composites_in_visibility_order(world, schedule):
    import torch.multiprocessing as mp
    npix, dims = 40 * 24, (64, 64, 64)
    # eyes outside the volume on either side of every axis: the front/back decision flips with the eye
    for eye in ((-80.0, -50.0, -70.0), (150.0, 140.0, 170.0), (-80.0, 140.0, -70.0)):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, npix, eye, dims, q, schedule)) for r in range(world)]
        for p in procs:
            p.start()
        got = q.get(timeout=120)
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        want = _expected(world, npix, np.array(eye), dims)
        assert np.abs(got - want).max() < 2e-6   # `over` is associative up to rounding
