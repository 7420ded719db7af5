"""TEST INFRASTRUCTURE ONLY - a probe, not a test: a KNOWN LIMITATION made visible.  Two ranks as host threads on the emulated
library (PDLP_B200_LIBRARY / PDLP_B200_NCCL_LIBRARY as in tests/test_full_emulation.py); the k-th device allocation of the
process fails.  Inside the matrix build the ranks agree on the failure and both return an error (k = 40); anywhere else the
failing rank returns its CUDA error and the OTHER RANK KEEPS WAITING in the next exchange (k = 3, 12, 25, 70, 110: "HUNG").
The remedy - abort the communicator and poison the peer flags from the failing rank - is not built.
    python oom_sharded_probe.py <k>"""
import ctypes, json, os, sys, threading
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/tests/emu_lib")
import numpy as np, scipy.sparse as sp
import fixtures, pdlp_b200
from pdlp_b200 import sharding
k = int(sys.argv[1])
countdown = ctypes.c_longlong.in_dll(pdlp_b200._lib, "_ZZ24pdlp_emu_fail_allocationE9countdown")
qp = fixtures.random_lp(40, 100, 4, 21)
p = pdlp_b200.PrimalDualHybridGradientParams(); p.termination_criteria.iteration_limit = 20
K = sp.csc_matrix(qp.constraint_matrix); world = 2
starts = sharding.column_partition(K.indptr, world); sharing = sharding.RowSharing.from_matrix(K, starts); uid = sharding.unique_id()
out = [None] * world
def rank_main(rank):
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    block = sharding.take_columns(qp, lo, hi, sharing.local_rows(rank))
    shard = sharding.ColumnShard(rank, world, int(starts[-1]), lo, uid, int(sharing.shared.size))
    try:
        r = pdlp_b200.primal_dual_hybrid_gradient(block, p, device=rank, shard=shard); out[rank] = "ok %d" % r.solve_log.iteration_count
    except Exception as e: out[rank] = repr(e)[:90]
countdown.value = k
ts = [threading.Thread(target=rank_main, args=(r,), daemon=True) for r in range(world)]
[t.start() for t in ts]; [t.join(40) for t in ts]
print(k, out, "HUNG" if any(t.is_alive() for t in ts) else "")
os._exit(0)
