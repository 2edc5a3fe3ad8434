"""A/B of the cross-shard exchange forms under torchrun, several configurations in ONE launch (start-up is what an
8-GPU call is charged for).  Each argument is form:share:scaling[:shape], e.g. rows:1.0:weak rows:0.8:weak ids:1.0:strong rows:1.0:weak:4.
Prints one JSON line per configuration (rank 0): the tick as bench.py times it (ShardRun.timed)."""
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rows = int(os.environ.get("AB_ROWS", "1000000000"))
    steps, warm = int(os.environ.get("AB_STEPS", "10")), 3
    for spec in sys.argv[1:]:
        form, share, scaling, *more = spec.split(":")
        share = float(share)
        if more:  # tiles per placement CTA of the fused push (TT_PUSH_VARIANT, read by the library at every tick)
            os.environ["TT_PUSH_VARIANT"] = more[0]
        else:
            os.environ.pop("TT_PUSH_VARIANT", None)
        total = rows * world if scaling == "weak" else rows
        root = int(total / world * share) // 8192 * 8192
        rest = total - root
        cut = lambda k: root + rest * k // (world - 1)
        first, last = (0, root) if rank == 0 else (cut(rank - 1), cut(rank))
        flags = {"rows": 0, "ids": tt.EXCHANGE_IDS}[form]
        run = bench.ShardRun(pkg, torch, dist, dev, local, rank, world, total, first, last - first, flags, local_ids=form != "ids")
        m = run.timed(steps, warm)
        every = [None] * world
        dist.all_gather_object(every, {"k1_ms": m["sweep_ms"], "k3g_ms": m["gather_ms"], "step_ms": m["step_ms_rank0"]})
        if rank == 0:
            assert m["collected"] == m["hits"], (m["collected"], m["hits"])
            print(json.dumps({"form": form, "push_shape": more[0] if more else "default", "root_share": share, "scaling": scaling, "n_gpus": world, "n_total": total,
                              "ms_per_step": m["ms_per_step"], "tasks_per_sec": total * steps / (m["total_ms"] * 1e-3),
                              "hits": m["hits"], "rank0": {"k1_ms": m["sweep_ms"], "k3g_ms": m["gather_ms"],
                                                           "collect_ms": m["collect_ms"], "step_ms": m["step_ms_rank0"]},
                              "peers_max": {k: max(e[k] for e in every[1:]) for k in ("k1_ms", "k3g_ms", "step_ms")}}), flush=True)
        run.close()
    dist.destroy_process_group()


main()
