"""Run under torchrun with N >= 2 GPUs (one rank per GPU, NCCL): checks both multi-GPU paths against a single-GPU render
of the same scene on rank 0 and prints one JSON line.

  sort-first  gathered frame must be BIT-IDENTICAL to the single-GPU frame of the same (ring, march_rq.cuh) kernel, and within 1e-3 of the
              queue kernel's frame (2^-31 fixed-point contributions, amplified by 1 / A at nearly transparent pixels)
  sort-last   N brick volumes (own voxels + ghosts), global sample lattice, binary swap over NCCL, un-premultiplied on rank 0:
              max-abs <= 1/255 and PSNR >= 45 dB against the single-GPU render of the whole volume

python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multigpu_check.py
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import ovr_b200
    from ovr_b200 import distributed as D
    capi = ovr_b200.capi
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = capi.lib()
    W, H, n = 640, 360, 160
    out = {"world": world}

    def renderer(scene, shading):
        r = ovr_b200.create_renderer("b200", local)
        r.shading = shading
        r.set_fbsize((W, H))
        r.init(scene, scene.camera)
        r.commit()
        r.use_stream(torch.cuda.current_stream().cuda_stream)   # ordered with the NCCL collectives
        return r

    # ---- sort-first (gradient + shadow shading, the queue kernel) ----
    sc = ovr_b200.scenes.noise_scene(n, tf="sparse", camera="oblique")
    r = renderer(sc, capi.SHADING_REF)
    if rank == 0:
        r.render()
        want_q = r.download(want_grad=False)[0].copy()     # undivided frame, queue kernel (float sums in march order)
        r.set_option("kernel", "rq4")                      # undivided frame, ring kernel: what the partitions must reproduce bit for bit
        r.render()
        want = r.download(want_grad=False)[0].copy()
        r.set_option("kernel", "auto")
    sf = D.SortFirst(r, rank, world)
    sf.step()
    torch.cuda.synchronize()
    if rank == 0:
        got = r.download(want_grad=False)[0]
        out["sort_first_bit_identical"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
        out["sort_first_vs_queue_kernel_max_abs"] = float(np.abs(got - want_q).max())
    # the same through the fused exchange: every rank stores its tiles straight into rank 0's frame
    dist.barrier()
    r._ck(L.b200vr_resize(r.handle, W + 64, H))   # reallocate: a fresh zeroed frame on every rank
    r._ck(L.b200vr_resize(r.handle, W, H))
    fp = D.SortFirstP2P(r, rank, world)
    fp.step()
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        got = r.download(want_grad=False)[0]
        out["sort_first_p2p_bit_identical"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
    fp.close()
    dist.barrier()
    r.close()

    # ---- sort-last ----
    full = ovr_b200.scenes.noise_scene(n, tf="ramp", camera="oblique")
    full.tfn.value_range = (0.0, 1.0)
    if rank == 0:
        ref = renderer(full, capi.SHADING_NONE)
        ref.render()
        want = ref.download(want_grad=False)[0].copy()
        ref.close()
    part, brick = D.brick_scene(full, rank, world)
    b = renderer(part, capi.SHADING_NONE)
    D.setup_brick_renderer(b, brick)
    b.render()
    swap = D.SortLast(rank, world, full.volume.dims, W * H, handle_renderer=b)
    img = swap.composite(D.frame_tensor(b), full.camera.frm)
    if rank == 0:
        b._ck(L.b200vr_finalize_premultiplied(b.handle, C.c_void_p(img.data_ptr()), W * H))
        torch.cuda.synchronize()
        got = img.cpu().numpy().reshape(H, W, 4)
        err = float(np.abs(got - want).max())
        mse = float(np.mean((got.astype(np.float64) - want) ** 2))
        out["sort_last_max_abs"] = err
        out["sort_last_psnr_db"] = float("inf") if mse == 0 else float(10 * np.log10(1.0 / mse))
        out["sort_last_ok"] = bool(err <= 1.0 / 255.0 and out["sort_last_psnr_db"] >= 45.0)
        out["composite_launches"] = swap.launches
    # the same frame through the fused peer-memory exchange must be bit-identical to the NCCL schedule
    b.render()
    p2p = D.SortLastP2P(rank, world, full.volume.dims, W * H, b)
    ptr = p2p.composite(full.camera.frm)
    if rank == 0:
        b._ck(L.b200vr_finalize_premultiplied(b.handle, C.c_void_p(ptr), W * H))
        torch.cuda.synchronize()
        got2 = b.download(want_grad=False)[0]
        out["sort_last_p2p_equals_nccl"] = bool(np.array_equal(got2.view(np.uint32), got.view(np.uint32)))
    # direct send, fused over peer memory (one kernel per rank, result straight into rank 0's other buffer), and over NCCL send/recv
    b.render()
    ptr = p2p.composite_direct(full.camera.frm)
    if rank == 0:
        torch.cuda.synchronize()
        got3 = torch.as_tensor(D._DevicePtr(ptr, W * H * 4), device="cuda").clone().cpu().numpy().reshape(H, W, 4)
        err3 = float(np.abs(got3 - want).max())
        mse3 = float(np.mean((got3.astype(np.float64) - want) ** 2))
        out["direct_send_p2p_max_abs"] = err3
        out["direct_send_p2p_psnr_db"] = float("inf") if mse3 == 0 else float(10 * np.log10(1.0 / mse3))
        out["direct_send_ok"] = bool(err3 <= 1.0 / 255.0 and out["direct_send_p2p_psnr_db"] >= 45.0)
    b.render()
    ds = D.SortLastDirect(rank, world, full.volume.dims, W * H, handle_renderer=b)
    img = ds.composite(D.frame_tensor(b), full.camera.frm)
    if rank == 0:
        b._ck(L.b200vr_finalize_premultiplied(b.handle, C.c_void_p(img.data_ptr()), W * H))
        torch.cuda.synchronize()
        got4 = img.cpu().numpy().reshape(H, W, 4)
        out["direct_send_p2p_equals_nccl"] = bool(np.array_equal(got4.view(np.uint32), got3.view(np.uint32)))
    dist.barrier()
    p2p.close()
    b.close()
    dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()
    if rank == 0 and not (out["sort_first_bit_identical"] and out["sort_first_p2p_bit_identical"] and out["sort_first_vs_queue_kernel_max_abs"] < 1e-3 and out["sort_last_ok"] and
                          out["sort_last_p2p_equals_nccl"] and out["direct_send_ok"] and out["direct_send_p2p_equals_nccl"]):
        sys.exit(1)


if __name__ == "__main__":
    main()
