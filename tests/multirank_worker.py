"""Worker of tests/test_multirank_gpu.py: one rank per GPU under torchrun. Runs the cell-sharded POD (resident and
row-streamed) and the halo-sharded projection through the library's own NCCL all-reduces and compares with the oracle on rank 0
(bench.parity_check: G 1e-13, eigenvalues 1e-10, modes 1e-8, operators 1e-10)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rank, local_rank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    env = bench.Env(rank, local_rank, world)
    pod = env.new_pod()
    res = bench.parity_check(env, pod)
    pod.close()
    ok = 1.0
    if rank == 0:
        print("MULTIRANK " + json.dumps(res), flush=True)
        ok = 1.0 if res["pass"] else 0.0
    ok = -env.max_over_ranks([-ok])[0]
    env.close()
    sys.exit(0 if ok > 0 else 1)


if __name__ == "__main__":
    main()
