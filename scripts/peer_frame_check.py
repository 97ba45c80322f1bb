"""Run under torchrun with N >= 2 GPUs: the frame assembled through the peer frame (NVLink stores into rank 0's
allocation) and through the NCCL gather must both equal the frame one GPU renders alone, byte for byte.
Usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/peer_frame_check.py [scene ...]"""
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("raytracing-project_b200")
sharding = importlib.import_module("raytracing-project_b200.sharding")


def main() -> None:
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rt.init(local)
    stream = torch.cuda.current_stream().cuda_stream
    scenes = sys.argv[1:] or ["scenes/snorlax_small.json", "scenes/synthetic_b_small.json", "scenes/penguin_small.json:paper"]
    bad = 0
    for tok in scenes:
        path, _, m = tok.partition(":")
        mode = rt.RenderMode.Paper if m == "paper" else rt.RenderMode.Standard
        scene = rt.load_scene_from_json(os.path.join(ROOT, path))
        W, H = scene.width, scene.height
        tracer = rt.Tracer(scene, mode)
        alone = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
        tracer.render_device(alone.data_ptr(), stream=stream, rank=0, world=1)
        part = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
        tracer.render_device(part.data_ptr(), stream=stream, rank=rank, world=world)
        gathered = sharding.gather_frame(part, W, H, int(mode), rank, world)
        peer = sharding.PeerFrame(rt, W, H, rank, world, dev)
        for _ in range(3):                      # re-rendering into the same frame must be idempotent
            tracer.render_device(peer.ptr, stream=stream, rank=rank, world=world)
            peer.complete()
        if rank == 0:
            a, b = bool(torch.equal(gathered, alone)), bool(torch.equal(peer.tensor(), alone))
            print(f"{tok}: {W}x{H} world={world} gather==alone {a} peer==alone {b}", flush=True)
            bad += (not a) + (not b)
        peer.close()
        scene.close()
    flag = torch.tensor([bad], device=dev)
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
