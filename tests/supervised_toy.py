"""Toy rank program for the supervisor tests: gloo all-reduce steps with a progress file as its "snapshot".
argv: <workdir> <steps> <mode>   mode: crash (rank 1 dies at step 3 of the first attempt), hang (rank 0 stops beating
in the first attempt), flaky (always dies while there is more than one rank), ok."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from veles_b200.parallel import supervisor


def main():
    workdir, steps, mode = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    attempt = supervisor.restart_count()
    snap = supervisor.resume_snapshot()
    start, total = 0, 0.0
    if snap:
        state = json.load(open(snap))
        start, total = state["step"], state["total"]
    for step in range(start, steps):
        supervisor.heartbeat()
        t = torch.tensor([float(step + 1)])
        dist.all_reduce(t)
        total += t.item() / world
        if mode == "crash" and attempt == 0 and rank == 1 and step == 3:
            os._exit(7)
        if mode == "hang" and attempt == 0 and rank == 0 and step == 2:
            time.sleep(3600)
        if mode == "flaky" and world > 1 and rank == world - 1 and step == start:
            os._exit(9)
        if rank == 0:
            tmp = os.path.join(workdir, "snap_%03d.json.tmp" % (step + 1))
            json.dump({"step": step + 1, "total": total, "world": world, "attempt": attempt}, open(tmp, "w"))
            os.replace(tmp, tmp[:-4])
        dist.barrier()
    if rank == 0:
        json.dump({"total": total, "world": world, "attempt": attempt, "resumed_from": snap},
                  open(os.path.join(workdir, "done.json"), "w"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
