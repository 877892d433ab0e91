#!/usr/bin/env python3
"""Where do the seconds go when N rank processes start on one box?  (multi-GPU CLI start-up)

Driver mode (no RANK in the environment): starts N copies of itself per variant and prints, per variant, the time from
spawn to each milestone (max over ranks): interpreter+import torch, CUDA context, process group rendezvous, first
all_reduce (communicator build), a gather-like batch of point-to-point sends to rank 0.
Variants: all GPUs visible to every rank / one GPU visible per rank (CUDA_VISIBLE_DEVICES), contexts created
before or after the rendezvous.
"""
import json
import os
import subprocess
import sys
import time


def rank_main():
    t = {"main": time.time()}
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev_index = int(os.environ["LOCAL_RANK"])
    order = os.environ.get("PROBE_ORDER", "ctx_first")
    import torch
    import torch.distributed as dist
    t["import"] = time.time()

    def ctx():
        torch.cuda.set_device(dev_index)
        torch.zeros(1, device="cuda").sum().item()
        t["ctx"] = time.time()

    def pg():
        dist.init_process_group("nccl", rank=rank, world_size=world)
        t["pg"] = time.time()

    if order == "ctx_first":
        ctx(); pg()
    else:
        pg(); ctx()
    x = torch.ones(1, device="cuda")
    dist.all_reduce(x)
    torch.cuda.synchronize()
    t["allreduce"] = time.time()
    assert int(x.item()) == world
    n = 1 << 18
    if rank == 0:
        bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(world - 1)]
        ops = [dist.P2POp(dist.irecv, b, r + 1) for r, b in enumerate(bufs)]
    else:
        ops = [dist.P2POp(dist.isend, torch.full((n,), rank, dtype=torch.uint8, device="cuda"), 0)]
    for r in dist.batch_isend_irecv(ops):
        r.wait()
    torch.cuda.synchronize()
    if rank == 0:
        assert all(int(b[0]) == r + 1 and int(b[-1]) == r + 1 for r, b in enumerate(bufs))
    t["p2p"] = time.time()
    dist.barrier()
    t["barrier"] = time.time()
    dist.destroy_process_group()
    t["destroy"] = time.time()
    print("PROBE " + json.dumps({"rank": rank, **t}), flush=True)


def driver():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    variants = [("all_visible,ctx_first", False, "ctx_first"), ("one_visible,ctx_first", True, "ctx_first"),
                ("all_visible,pg_first", False, "pg_first"), ("one_visible,pg_first,nvls_off", True, "pg_first")]
    for k, (name, isolate, order) in enumerate(variants):
        base = dict(os.environ, WORLD_SIZE=str(n), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29650 + k), PROBE_ORDER=order)
        if "nvls_off" in name:
            base["NCCL_NVLS_ENABLE"] = "0"
        t0 = time.time()
        kids = []
        for r in range(n):
            env = dict(base, RANK=str(r), LOCAL_RANK="0" if isolate else str(r))
            if isolate:
                env["CUDA_VISIBLE_DEVICES"] = str(r)
            kids.append(subprocess.Popen([sys.executable, __file__], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
        rows, ok = [], True
        for kproc in kids:
            try:
                out, err = kproc.communicate(timeout=150)
            except subprocess.TimeoutExpired:
                kproc.kill()
                out, err = kproc.communicate()
                ok = False
            if kproc.returncode != 0:
                ok = False
                sys.stderr.write(err[-1500:] + "\n")
            rows += [json.loads(l[6:]) for l in out.splitlines() if l.startswith("PROBE ")]
        wall = time.time() - t0
        res = {"variant": name, "ranks": n, "ok": ok, "wall_s": round(wall, 2)}
        if rows:
            prev = None
            for key in ("main", "import", "ctx", "pg", "allreduce", "p2p", "barrier", "destroy"):
                if all(key in r for r in rows):
                    res[key] = {"max": round(max(r[key] for r in rows) - t0, 2), "rank0": round([r[key] for r in rows if r["rank"] == 0][0] - t0, 2)}
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    if "RANK" in os.environ:
        rank_main()
    else:
        driver()
