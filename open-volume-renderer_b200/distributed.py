"""Multi-GPU rendering, one process per GPU (the reference itself is single-device, device_impl.cpp:370-372).

* sort-first (the volume fits on every GPU): every rank holds the whole scene and renders the 64x32-pixel super tiles it owns
  (round robin, csrc/launch_params.cuh::warp_tile_origin); ONE exchange step gathers the dense tile slices on the display rank.
  The assembled frame is bit-identical to the single-GPU frame (pixels are independent).
* sort-last (the volume does not fit): the volume is cut into a power-of-two grid of bricks, one per rank. Every rank marches
  ALL pixels against its brick on the GLOBAL sample lattice (b200vr_set_lattice_box) into a premultiplied RGBA image; the
  images are merged front to back by binary swap (log2 N rounds, each exchanging half of the rank's current pixel range
  with one partner) and the N final strips are gathered on the display rank, which un-premultiplies once
  (alpha_blend, shaders_common.h:329-337). Ray marching with SHADING_NONE only: shadow rays and random walks leave a brick.

The exchange runs over torch.distributed (NCCL on the GPUs, over NVLink / NVSwitch); the blend is the CUDA kernel behind
b200vr_composite_over. The schedule itself (who talks to whom, which half, who is in front) is plain integer logic and is
unit-tested on CPU with the gloo backend and an injected blend function; without such an injection this module only accepts
CUDA tensors -- there is no CPU rendering path.
"""
import ctypes as C

import numpy as np

SUPER_TILE_W, SUPER_TILE_H = 64, 32   # csrc/launch_params.cuh


# ----------------------------------------------------------------------------------------------------------------------
# sort-first
# ----------------------------------------------------------------------------------------------------------------------
def super_tile_grid(width, height):
    return (width + SUPER_TILE_W - 1) // SUPER_TILE_W, (height + SUPER_TILE_H - 1) // SUPER_TILE_H


def owned_super_tiles(width, height, rank, nranks):
    """Indices (row-major over the super-tile grid) of the tiles `rank` renders: m * nranks + rank."""
    sw, sh = super_tile_grid(width, height)
    return list(range(rank, sw * sh, nranks))


def tile_slice_floats(width, height, rank, nranks):
    """Size of the dense rgba slice b200vr_pack_tiles writes for `rank` (float count)."""
    return len(owned_super_tiles(width, height, rank, nranks)) * SUPER_TILE_W * SUPER_TILE_H * 4


def assemble_slices(slices, width, height):
    """Host restatement of b200vr_unpack_tiles for tests: slices[rank] is a flat float32 array laid out
    [owned tile m][y in 32][x in 64][rgba]; returns the [H, W, 4] frame."""
    nranks = len(slices)
    sw, _ = super_tile_grid(width, height)
    frame = np.zeros((height, width, 4), np.float32)
    for rank, sl in enumerate(slices):
        tiles = owned_super_tiles(width, height, rank, nranks)
        sl = np.asarray(sl, np.float32).reshape(len(tiles), SUPER_TILE_H, SUPER_TILE_W, 4)
        for m, t in enumerate(tiles):
            x0, y0 = (t % sw) * SUPER_TILE_W, (t // sw) * SUPER_TILE_H
            w, h = min(SUPER_TILE_W, width - x0), min(SUPER_TILE_H, height - y0)
            frame[y0:y0 + h, x0:x0 + w] = sl[m, :h, :w]
    return frame


class SortFirst:
    """Renders this rank's tiles and gathers the frame on `dst`. `renderer` is an ovr_b200 MainRenderer whose scene is
    already committed on every rank."""

    def __init__(self, renderer, rank, world, dst=0, group=None):
        import torch
        from . import _capi as capi
        self.r, self.rank, self.world, self.dst, self.group = renderer, rank, world, dst, group
        self.L = capi.lib()
        renderer._ck(self.L.b200vr_set_tile_partition(renderer.handle, rank, world))
        n = max(int(self.L.b200vr_tile_slice_floats(renderer.handle, k, world)) for k in range(world))
        self.slice = torch.zeros(n, dtype=torch.float32, device="cuda")
        self.gathered = ([torch.zeros(n, dtype=torch.float32, device="cuda") for _ in range(world)]
                         if world > 1 and rank == dst else None)
        self.launches = 0

    def step(self):
        """One frame: render own tiles, pack, gather, scatter the other ranks' slices into the display rank's frame."""
        import torch.distributed as dist
        r, L = self.r, self.L
        r.render()
        self.launches += r.last_stats["kernel_launches"]
        if self.world > 1:
            r._ck(L.b200vr_pack_tiles(r.handle, C.c_void_p(self.slice.data_ptr())))
            self.launches += 1
            dist.gather(self.slice, self.gathered, dst=self.dst, group=self.group)
            if self.rank == self.dst:
                for k in range(self.world):
                    if k != self.rank:
                        r._ck(L.b200vr_unpack_tiles(r.handle, C.c_void_p(self.gathered[k].data_ptr()), k, self.world))
                        self.launches += 1
        return r.last_stats


# ----------------------------------------------------------------------------------------------------------------------
# sort-last: brick decomposition
# ----------------------------------------------------------------------------------------------------------------------
GHOST_LO, GHOST_HI = 1, 2   # voxels a brick keeps beyond its cut planes: tri-linear footprint + forward-difference gradient


def brick_grid(nranks):
    """Power-of-two rank counts -> bricks per axis (bx, by, bz), split x first: 1 -> (1,1,1), 2 -> (2,1,1), 4 -> (2,2,1),
    8 -> (2,2,2), 16 -> (4,2,2) ... Bit b of a rank selects the half along axis b % 3 at level b // 3."""
    if nranks < 1 or nranks & (nranks - 1):
        raise ValueError("sort-last needs a power-of-two number of ranks")
    g = [1, 1, 1]
    for b in range(nranks.bit_length() - 1):
        g[b % 3] *= 2
    return tuple(g)


def brick_index(rank, nranks):
    """(i, j, k) of the brick of `rank`: the bits of the rank, dealt to the axes round robin (bit b -> axis b % 3)."""
    idx = [0, 0, 0]
    for b in range(nranks.bit_length() - 1):
        idx[b % 3] |= ((rank >> b) & 1) << (b // 3)
    return tuple(idx)


class Brick:
    """Voxel and object-space extents of one brick of a dims = (nx, ny, nz) volume.

    own_lo/own_hi     voxel range [lo, hi) the brick is responsible for (cut at multiples of dims / grid)
    org, size         voxel range it stores: own range grown by the ghost layers, clamped to the volume
    clip_lo/clip_hi   the own range in the BRICK's object space (unit cube = stored voxels)
    own_lo_w/own_hi_w the own range in world space                                            -> b200vr_set_brick
    vol_lo_w/vol_hi_w the whole volume in world space (same floats on every rank)             -> b200vr_set_brick
    obj_to_world      3x4 row-major: brick unit cube -> world, such that all bricks tile the world box of the whole volume
    """

    def __init__(self, dims, rank, nranks, grid_spacing=(1.0, 1.0, 1.0), grid_origin=(0.0, 0.0, 0.0)):
        g = brick_grid(nranks)
        idx = brick_index(rank, nranks)
        self.grid, self.index = g, idx
        self.own_lo = tuple(dims[a] * idx[a] // g[a] for a in range(3))
        self.own_hi = tuple(dims[a] * (idx[a] + 1) // g[a] for a in range(3))
        self.org = tuple(max(self.own_lo[a] - GHOST_LO, 0) for a in range(3))
        end = tuple(min(self.own_hi[a] + GHOST_HI, dims[a]) for a in range(3))
        self.size = tuple(end[a] - self.org[a] for a in range(3))
        f = np.float32
        self.clip_lo = tuple(f(self.own_lo[a] - self.org[a]) / f(self.size[a]) for a in range(3))
        self.clip_hi = tuple(f(self.own_hi[a] - self.org[a]) / f(self.size[a]) for a in range(3))
        # world-space planes: a pure function of (origin, spacing, voxel index), hence identical floats on all ranks
        w = lambda a, v: f(f(grid_origin[a]) + f(grid_spacing[a]) * f(v))   # noqa: E731
        self.own_lo_w = tuple(w(a, self.own_lo[a]) for a in range(3))
        self.own_hi_w = tuple(w(a, self.own_hi[a]) for a in range(3))
        self.vol_lo_w = tuple(w(a, 0) for a in range(3))
        self.vol_hi_w = tuple(w(a, dims[a]) for a in range(3))
        # whole volume: otw = translate(grid_origin) * scale(grid_spacing * dims) (device_impl.cpp:288-296)
        m = np.zeros((3, 4), np.float32)
        for a in range(3):
            m[a, a] = f(grid_spacing[a]) * f(self.size[a])
            m[a, 3] = f(grid_origin[a]) + f(grid_spacing[a]) * f(self.org[a])
        self.obj_to_world = m
        # cut planes in world space per axis level, for the visibility order
        self.dims, self.spacing, self.origin = tuple(dims), tuple(grid_spacing), tuple(grid_origin)


def split_plane(dims, nranks, bit, rank, grid_spacing=(1.0, 1.0, 1.0), grid_origin=(0.0, 0.0, 0.0)):
    """World coordinate (axis, value) of the plane that separates `rank` from `rank ^ (1 << bit)`.
    At round `bit` of the swap both partners hold the composite of all bricks that differ from them only in bits < bit,
    a convex box cut from the partner's box by exactly this plane."""
    axis, level = bit % 3, bit // 3
    g = brick_grid(nranks)[axis]
    idx = brick_index(rank, nranks)[axis]
    # partner differs in bit `level` of the axis index: the plane is the upper face of the lower of the two bricks at
    # the granularity of that bit
    lo_idx = (idx >> level) & ~1
    plane_vox = dims[axis] * (((lo_idx + 1) << level)) // g
    return axis, float(grid_origin[axis]) + float(grid_spacing[axis]) * float(plane_vox)


def in_front(eye, dims, nranks, bit, rank, **kw):
    """True when `rank`'s side of the round-`bit` plane is the eye's side: every ray then leaves `rank`'s box before it
    enters the partner's (two convex sets separated by a plane, rays start at the eye), for ALL pixels at once."""
    axis, plane = split_plane(dims, nranks, bit, rank, **kw)
    level = bit // 3
    mine_low = ((brick_index(rank, nranks)[axis] >> level) & 1) == 0
    eye_low = float(eye[axis]) < plane
    return mine_low == eye_low


# ----------------------------------------------------------------------------------------------------------------------
# sort-last: binary swap
# ----------------------------------------------------------------------------------------------------------------------
def binary_swap_plan(rank, nranks, n_pixels, align=1):
    """[(partner, keep (lo, hi), send (lo, hi))] per round and the final owned range. Round b pairs rank with rank ^ (1 << b);
    the current range is halved (at a multiple of `align` pixels), the rank whose bit b is 0 keeps the lower half."""
    lo, hi = 0, n_pixels
    plan = []
    for b in range(nranks.bit_length() - 1):
        mid = lo + ((hi - lo) // 2 // align) * align
        partner = rank ^ (1 << b)
        if (rank >> b) & 1:
            keep, send = (mid, hi), (lo, mid)
        else:
            keep, send = (lo, mid), (mid, hi)
        plan.append((partner, keep, send))
        lo, hi = keep
    return plan, (lo, hi)


def final_ranges(nranks, n_pixels, align=1):
    return [binary_swap_plan(r, nranks, n_pixels, align)[1] for r in range(nranks)]


def over_torch(front, back):
    """front OVER back on premultiplied [n, 4] pixels -- the CPU stand-in the gloo tests inject; the product path blends
    with b200vr_composite_over."""
    return front + (1.0 - front[:, 3:4]) * back


class SortLast:
    """Binary-swap compositing of per-rank premultiplied images.

    image     flat float32 tensor of n_pixels * 4 on this rank (CUDA in production)
    eye       camera position in world space (decides front/back per round)
    blend     None -> b200vr_composite_over through `handle_renderer`; tests pass `over_torch`
    """

    def __init__(self, rank, world, dims, n_pixels, handle_renderer=None, blend=None, group=None, dst=0,
                 grid_spacing=(1.0, 1.0, 1.0), grid_origin=(0.0, 0.0, 0.0), align=32):
        self.rank, self.world, self.dims, self.n = rank, world, tuple(dims), n_pixels
        self.r, self.blend, self.group, self.dst = handle_renderer, blend, group, dst
        self.kw = dict(grid_spacing=grid_spacing, grid_origin=grid_origin)
        self.plan, self.mine = binary_swap_plan(rank, world, n_pixels, align)
        self.ranges = final_ranges(world, n_pixels, align)
        self.recv = None
        self.launches = 0

    def _over(self, front, back, out):
        if self.blend is not None:
            out.copy_(self.blend(front.view(-1, 4), back.view(-1, 4)).view(-1))
            return
        if not front.is_cuda:
            raise RuntimeError("SortLast: CPU tensors need an injected blend function; the product path is CUDA only")
        from . import _capi as capi
        self.r._ck(capi.lib().b200vr_composite_over(self.r.handle, C.c_void_p(front.data_ptr()), C.c_void_p(back.data_ptr()),
                                                    C.c_void_p(out.data_ptr()), front.numel() // 4))
        self.launches += 1

    def composite(self, image, eye):
        """In place on `image`; returns the [n_pixels * 4] final premultiplied frame on `dst` (None elsewhere)."""
        import torch
        import torch.distributed as dist
        if self.recv is None or self.recv.numel() < self.n * 2 or self.recv.device != image.device:
            self.recv = torch.empty(self.n * 2, dtype=torch.float32, device=image.device)   # largest half: n/2 pixels
        for b, (partner, keep, send) in enumerate(self.plan):
            k0, k1 = keep[0] * 4, keep[1] * 4
            s0, s1 = send[0] * 4, send[1] * 4
            got = self.recv[:k1 - k0]
            ops = [dist.P2POp(dist.isend, image[s0:s1], partner, self.group), dist.P2POp(dist.irecv, got, partner, self.group)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            mine = image[k0:k1]
            if in_front(eye, self.dims, self.world, b, self.rank, **self.kw):
                self._over(mine, got, mine)
            else:
                self._over(got, mine, mine)
        # gather the N strips (equal length up to rounding: pad to the longest)
        if self.world == 1:
            return image
        longest = max(hi - lo for lo, hi in self.ranges) * 4
        strip = torch.zeros(longest, dtype=torch.float32, device=image.device)
        lo, hi = self.mine
        strip[:(hi - lo) * 4] = image[lo * 4:hi * 4]
        parts = [torch.empty_like(strip) for _ in range(self.world)] if self.rank == self.dst else None
        dist.gather(strip, parts, dst=self.dst, group=self.group)
        if self.rank != self.dst:
            return None
        for k, (lo, hi) in enumerate(self.ranges):
            image[lo * 4:hi * 4] = parts[k][:(hi - lo) * 4]
        return image


# ----------------------------------------------------------------------------------------------------------------------
# sort-last: direct send
# ----------------------------------------------------------------------------------------------------------------------
def visibility_order(eye, dims, nranks, **kw):
    """Ranks front to back as seen from `eye`, for ALL pixels at once: the brick grid is a BSP tree (bit b of a rank picks
    the side of the level-b plane), so of two ranks the one on the eye's side of the plane of their HIGHEST differing bit
    is in front -- the same planes, in the same order, that the rounds of the binary swap use."""
    import functools

    def cmp(a, b):
        hb = (a ^ b).bit_length() - 1
        return -1 if in_front(eye, dims, nranks, hb, a, **kw) else 1

    return sorted(range(nranks), key=functools.cmp_to_key(cmp))


def direct_send_ranges(nranks, n_pixels, align=32):
    """Pixel range [lo, hi) rank k composites under direct send: n_pixels cut into nranks nearly equal strips at multiples
    of `align` pixels (the last strip takes the remainder)."""
    per = ((n_pixels + nranks - 1) // nranks + align - 1) // align * align
    return [(min(k * per, n_pixels), min((k + 1) * per, n_pixels)) for k in range(nranks)]


class SortLastDirect:
    """Direct-send compositing over send/recv (NCCL on the GPUs; gloo with an injected blend in the CPU tests): every rank
    owns one strip of the image, receives that strip of every other rank's layer, blends the N strips front to back in
    visibility order and the display rank gathers the N results. One exchange round instead of log2 N, N - 1 messages per
    rank instead of log2 N. The production path is SortLastDirectP2P; this one is its cross-check."""

    def __init__(self, rank, world, dims, n_pixels, handle_renderer=None, blend=None, group=None, dst=0,
                 grid_spacing=(1.0, 1.0, 1.0), grid_origin=(0.0, 0.0, 0.0), align=32):
        self.rank, self.world, self.dims, self.n = rank, world, tuple(dims), n_pixels
        self.r, self.blend, self.group, self.dst = handle_renderer, blend, group, dst
        self.kw = dict(grid_spacing=grid_spacing, grid_origin=grid_origin)
        self.ranges = direct_send_ranges(world, n_pixels, align)
        self.mine = self.ranges[rank]
        self.launches = 0

    def composite(self, image, eye):
        """Returns the [n_pixels * 4] final premultiplied frame on `dst` (in place on `image`), None elsewhere."""
        import torch
        import torch.distributed as dist
        lo, hi = self.mine
        n = (hi - lo) * 4
        strips = {self.rank: image[lo * 4:hi * 4].clone()}
        ops = []
        for k in range(self.world):
            if k == self.rank:
                continue
            klo, khi = self.ranges[k]
            strips[k] = torch.empty(n, dtype=torch.float32, device=image.device)
            if khi > klo:
                ops.append(dist.P2POp(dist.isend, image[klo * 4:khi * 4], k, self.group))
            if n > 0:
                ops.append(dist.P2POp(dist.irecv, strips[k], k, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        order = visibility_order(eye, self.dims, self.world, **self.kw)
        if n > 0:
            if self.blend is not None:
                acc = strips[order[0]].view(-1, 4)
                for k in order[1:]:
                    acc = self.blend(acc, strips[k].view(-1, 4))
                out = acc.reshape(-1)
            else:
                if not image.is_cuda:
                    raise RuntimeError("SortLastDirect: CPU tensors need an injected blend function; the product path is CUDA only")
                from . import _capi as capi
                out = torch.empty(n, dtype=torch.float32, device=image.device)
                ptrs = (C.c_void_p * self.world)(*[strips[k].data_ptr() for k in order])
                self.r._ck(capi.lib().b200vr_composite_layers(self.r.handle, ptrs, self.world, C.c_void_p(out.data_ptr()), hi - lo, 0))
                self.launches += 1
        else:
            out = torch.empty(0, dtype=torch.float32, device=image.device)
        if self.world == 1:
            image[lo * 4:hi * 4] = out
            return image
        longest = max(b - a for a, b in self.ranges) * 4
        strip = torch.zeros(longest, dtype=torch.float32, device=image.device)
        strip[:n] = out
        parts = [torch.empty_like(strip) for _ in range(self.world)] if self.rank == self.dst else None
        dist.gather(strip, parts, dst=self.dst, group=self.group)
        if self.rank != self.dst:
            return None
        for k, (a, b) in enumerate(self.ranges):
            image[a * 4:b * 4] = parts[k][:(b - a) * 4]
        return image


# ----------------------------------------------------------------------------------------------------------------------
# sort-last: per-rank scene + renderer setup
# ----------------------------------------------------------------------------------------------------------------------
def brick_scene(scene, rank, nranks):
    """(Scene holding only this rank's brick, Brick). The brick keeps the camera, TF and sampling rate of `scene`; its
    volume is the stored voxel range (own range + ghosts), placed in world space so that the bricks tile the volume."""
    from .scenes import Scene, Volume
    vol = scene.volume
    if not scene.tfn.value_range[0] <= scene.tfn.value_range[1]:
        raise ValueError("sort-last needs an explicit scalar mapping range: a brick's own data range is not the volume's")
    b = Brick(vol.dims, rank, nranks, vol.grid_spacing, vol.grid_origin)
    origin = tuple(float(np.float32(vol.grid_origin[a]) + np.float32(vol.grid_spacing[a]) * np.float32(b.org[a])) for a in range(3))
    if vol.data is not None:
        (x0, y0, z0), (sx, sy, sz) = b.org, b.size
        part = Volume(data=np.ascontiguousarray(vol.data[z0:z0 + sz, y0:y0 + sy, x0:x0 + sx]), grid_origin=origin,
                      grid_spacing=vol.grid_spacing)
    else:
        org0 = vol.noise["org"]
        part = Volume(noise={"seed": vol.noise["seed"], "org": tuple(int(org0[a]) + b.org[a] for a in range(3)), "dims": b.size},
                      grid_origin=origin, grid_spacing=vol.grid_spacing)
    return Scene(part, scene.tfn, scene.camera, scene.volume_sampling_rate), b


def setup_brick_renderer(renderer, brick):
    """After init(): premultiplied output, no shading, world-space own box + whole-volume box (b200vr_set_brick)."""
    from . import _capi as capi
    L = capi.lib()
    renderer.shading = capi.SHADING_NONE
    renderer.premultiplied = True
    renderer.apply_device_options()
    f3 = C.c_float * 3
    v = lambda t: f3(*[float(x) for x in t])   # noqa: E731
    renderer._ck(L.b200vr_set_brick(renderer.handle, v(brick.own_lo_w), v(brick.own_hi_w), v(brick.vol_lo_w), v(brick.vol_hi_w)))


class _DevicePtr:
    """Zero-copy torch view of a raw device pointer (torch.as_tensor reads __cuda_array_interface__)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (int(n_floats),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def frame_tensor(renderer):
    """The renderer's current rgba buffer (W * H * 4 floats) as a CUDA tensor, without a copy; valid until the next
    resize / swap on that renderer."""
    import torch
    fb = renderer.mapframe()
    return torch.as_tensor(_DevicePtr(fb.rgba[0], fb.rgba[1] // 4), device="cuda")


class SortLastP2P:
    """Binary swap with the exchange FUSED into the blend: every rank maps its peers' frame buffers (CUDA IPC, NVLink peer
    access) and `b200vr_composite_over` reads the partner's half image straight out of the partner's memory while it blends
    -- no send/recv, no staging buffer. Round b: rank keeps one half of its current pixel range and writes
    keep = front OVER back there, `back` or `front` being the partner's pixels of the same range; the partner does the
    same on the other half, so nobody writes what somebody else reads within a round. Rounds, the final pull of the N
    strips by the display rank and the next frame's render are separated by barriers (stream-ordered: the NCCL barrier is
    enqueued behind the blends on each rank's stream). All calls are on the caller's stream (b200vr_set_stream)."""

    def __init__(self, rank, world, dims, n_pixels, renderer, group=None, dst=0, grid_spacing=(1.0, 1.0, 1.0),
                 grid_origin=(0.0, 0.0, 0.0), align=32):
        import torch.distributed as dist
        from . import _capi as capi
        self.rank, self.world, self.dims, self.n = rank, world, tuple(dims), n_pixels
        self.r, self.group, self.dst = renderer, group, dst
        self.L = capi.lib()
        self.kw = dict(grid_spacing=grid_spacing, grid_origin=grid_origin)
        self.plan, self.mine = binary_swap_plan(rank, world, n_pixels, align)
        self.ranges = final_ranges(world, n_pixels, align)
        self.direct_ranges = direct_send_ranges(world, n_pixels, align)
        self.launches = 0
        # exchange the IPC handles of both frame buffers, open everybody else's
        mine = []
        for b in (0, 1):
            buf = (C.c_ubyte * 64)()
            renderer._ck(self.L.b200vr_ipc_export_frame(renderer.handle, b, buf))
            mine.append(bytes(buf))
        everyone = [None] * world
        dist.all_gather_object(everyone, mine, group=group)
        self._tok = None
        self.peer = {}          # (rank, buffer) -> device pointer in THIS process
        for k in range(world):
            if k == rank:
                continue
            for b in (0, 1):
                p = C.c_void_p()
                h = (C.c_ubyte * 64).from_buffer_copy(everyone[k][b])
                renderer._ck(self.L.b200vr_ipc_open(renderer.handle, h, C.byref(p)))
                self.peer[(k, b)] = p.value

    def _barrier(self):
        """All ranks' streams have reached this point: a one-word all-reduce enqueued on the current stream (the host does
        not wait; dist.barrier() blocks the host and cost ~0.2 ms per call here)."""
        import torch
        import torch.distributed as dist
        if self._tok is None:
            self._tok = torch.zeros(1, dtype=torch.float32, device="cuda")
        dist.all_reduce(self._tok, group=self.group)

    def close(self):
        for p in self.peer.values():
            self.L.b200vr_ipc_close(self.r.handle, C.c_void_p(p))
        self.peer = {}

    def composite(self, eye):
        """In place on the renderer's current frame buffer; after the call rank `dst` holds the premultiplied frame."""
        import torch.distributed as dist
        r, L = self.r, self.L
        cur = int(L.b200vr_current_buffer(r.handle))
        own = r.mapframe().rgba[0]
        for b, (partner, keep, _send) in enumerate(self.plan):
            self._barrier()   # the partner's pixels of this round are final (render / previous round)
            off, n = keep[0] * 16, keep[1] - keep[0]
            mine, theirs = own + off, self.peer[(partner, cur)] + off
            front, back = (mine, theirs) if in_front(eye, self.dims, self.world, b, self.rank, **self.kw) else (theirs, mine)
            r._ck(L.b200vr_composite_over(r.handle, C.c_void_p(front), C.c_void_p(back), C.c_void_p(mine), n))
            self.launches += 1
        self._barrier()
        if self.rank == self.dst:            # pull the other ranks' strips over NVLink
            for k, (lo, hi) in enumerate(self.ranges):
                if k != self.rank and hi > lo:
                    r._ck(L.b200vr_copy_device(r.handle, C.c_void_p(own + lo * 16), C.c_void_p(self.peer[(k, cur)] + lo * 16), (hi - lo) * 16))
                    self.launches += 1
        self._barrier()       # nobody renders the next frame into a buffer that is still being read
        return own if self.rank == self.dst else None

    def composite_direct(self, eye, finalize=True):
        """Direct send FUSED with the blend over NVLink peer memory: after one barrier every rank runs ONE kernel
        (b200vr_composite_layers) over the strip it owns -- it reads that strip of all N layers straight out of the GPUs that
        rendered them, blends them front to back in visibility order, un-premultiplies and stores the result straight into
        the display rank's OTHER frame buffer (the one nobody rendered into this frame) -- then one more barrier. Two
        barriers and one launch per rank and frame against log2 N + 2 barriers, log2 N blends and N - 1 pulls for the binary
        swap; the bytes over NVLink are the same (N - 1) / N of a frame per rank. Returns, on the display rank, the device
        pointer of the final frame (its buffer `1 - current`), None elsewhere."""
        r, L = self.r, self.L
        cur = int(L.b200vr_current_buffer(r.handle))
        own = r.mapframe().rgba[0]
        lo, hi = self.direct_ranges[self.rank]
        order = visibility_order(eye, self.dims, self.world, **self.kw)
        base = {k: (own if k == self.rank else self.peer[(k, cur)]) for k in range(self.world)}
        if self.rank == self.dst:
            p = C.c_void_p()
            r._ck(L.b200vr_map_index(r.handle, cur ^ 1, C.byref(p), None))
            out_base = p.value
        else:
            out_base = self.peer[(self.dst, cur ^ 1)]
        self._barrier()   # every rank's layer is rendered
        if hi > lo:
            ptrs = (C.c_void_p * self.world)(*[base[k] + lo * 16 for k in order])
            r._ck(L.b200vr_composite_layers(r.handle, ptrs, self.world, C.c_void_p(out_base + lo * 16), hi - lo, 1 if finalize else 0))
            self.launches += 1
        self._barrier()   # all strips have landed in the display rank's buffer; the layers may be overwritten by the next frame
        return out_base if self.rank == self.dst else None


class SortLastDirectP2P(SortLastP2P):
    """The direct-send schedule as its own class: composite() = SortLastP2P.composite_direct()."""

    def composite(self, eye, finalize=True):
        return self.composite_direct(eye, finalize)


class SortFirstP2P:
    """Sort-first with the exchange fused into the render kernel: every rank opens the display rank's frame buffers (CUDA
    IPC, NVLink peer access) and renders its tiles STRAIGHT INTO THEM (b200vr_set_frame_target) -- the kernel's pixel stores
    are the transfer; there is no pack / gather / unpack. One stream-ordered one-word all-reduce per frame tells the
    display rank that all tiles have landed. Ranks must call swap() in lockstep (they share the display rank's buffer index)."""

    def __init__(self, renderer, rank, world, dst=0, group=None):
        import torch
        import torch.distributed as dist
        from . import _capi as capi
        self.r, self.rank, self.world, self.dst, self.group = renderer, rank, world, dst, group
        self.L = capi.lib()
        renderer._ck(self.L.b200vr_set_tile_partition(renderer.handle, rank, world))
        self.launches = 0
        self._tok = torch.zeros(1, dtype=torch.float32, device="cuda") if world > 1 else None
        self.remote = {}
        if world > 1:
            handles = [None]
            if rank == dst:
                hs = []
                for b in range(4):
                    buf = (C.c_ubyte * 64)()
                    renderer._ck(self.L.b200vr_ipc_export_frame(renderer.handle, b, buf))
                    hs.append(bytes(buf))
                handles = [hs]
            dist.broadcast_object_list(handles, src=dst, group=group)
            if rank != dst:
                for b in range(4):
                    p = C.c_void_p()
                    h = (C.c_ubyte * 64).from_buffer_copy(handles[0][b])
                    renderer._ck(self.L.b200vr_ipc_open(renderer.handle, h, C.byref(p)))
                    self.remote[b] = p.value

    def close(self):
        if self.remote:
            self.r._ck(self.L.b200vr_set_frame_target(self.r.handle, None, None))
            for p in self.remote.values():
                self.L.b200vr_ipc_close(self.r.handle, C.c_void_p(p))
            self.remote = {}

    def step(self, wait=True):
        """One frame. wait=False: nothing here blocks the host -- the frame and the completion all-reduce are enqueued on the
        stream (b200vr_render_async) and the caller collects the counters of all frames since with renderer.wait()."""
        import torch.distributed as dist
        r, L = self.r, self.L
        if self.remote:
            cur = int(L.b200vr_current_buffer(r.handle))
            r._ck(L.b200vr_set_frame_target(r.handle, C.c_void_p(self.remote[cur]), C.c_void_p(self.remote[2 + cur])))
        if wait:
            r.render()
            self.launches += r.last_stats["kernel_launches"]
        else:
            r.render_async()
        if self.world > 1:
            dist.all_reduce(self._tok, group=self.group)   # stream-ordered: every rank's tiles are in the display rank's frame
        return r.last_stats if wait else None
