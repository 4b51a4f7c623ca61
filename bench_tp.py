"""Multi-GPU leg of bench.py (filled in with the tensor-parallel path; see cpmcu_b200/sharding.py)."""
import json


def main_tp(args, rank, world, local_rank):
    if rank == 0:
        print(json.dumps({"metric": "w4a16_gemm_weight_stream_GBps", "n_gpus": world,
                          "unavailable": "tensor-parallel leg not built yet"}), flush=True)
