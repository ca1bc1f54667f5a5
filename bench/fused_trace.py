"""Phase timestamps of the fused kernels (2+ GPUs, run under torchrun inside `timeout`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from b200dist.launch import env as launch_env
info = launch_env.init_distributed("nccl")
rank, world = info.rank, info.world_size
dev = torch.device("cuda", torch.cuda.current_device())
from b200dist.parallel import fabric as fb
from b200dist.ops import native as N
F = fb.get_fabric(dist.group.WORLD)
L = N.lib()
dbg = torch.zeros(148 * 8, dtype=torch.int64, device=dev)

def show(tag, ncomm=0):
    d = dbg.view(148, 8).cpu()
    t0 = d[:, 0][d[:, 0] > 0].min().item()
    rel = lambda col, rows: [(x - t0) / 1e3 for x in d[rows, col].tolist() if x > 0]
    import statistics as st
    def summ(col, rows):
        v = rel(col, rows)
        return f"n={len(v)} min={min(v):.1f} med={st.median(v):.1f} max={max(v):.1f}" if v else "-"
    if rank == 0:
        allr = list(range(148))
        print(f"--- {tag} (us since first CTA start)")
        if ncomm:
            cr, gr = list(range(ncomm)), list(range(ncomm, 148))
            print("  comm CTAs: start", summ(0, cr), "| first peer ready", summ(1, cr), "| pulls done", summ(2, cr))
            print("  gemm CTAs: start", summ(0, gr), "| tiles done", summ(1, gr), "| exit", summ(4, gr))
        else:
            print("  start", summ(0, allr), "| owed signal sent before own tile", summ(7, allr), "| own-tile wait begins", summ(3, allr), "| own-tile wait ends", summ(6, allr))
            print("  tile loop done", summ(1, allr), "| last signal sent", summ(2, allr), "| CTA exit", summ(4, allr))
        print("  kernel complete (last CTA)", summ(5, allr))

def trace_ag(rows, K, Nn, tag):
    x = F.shard_buffer(rows, K, torch.bfloat16); x.normal_()
    W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
    L.b2_fused_set_debug(None)
    for i in range(3):
        F.ag_gemm(x, W)
    torch.cuda.synchronize(); dist.barrier()
    L.b2_fused_set_debug(N.c_p(dbg.data_ptr()))
    dbg.zero_(); F.ag_gemm(x, W); torch.cuda.synchronize(); dist.barrier()
    show(f"ag_gemm {tag} rows/rank {rows} K{K} N{Nn}", ncomm=int(os.environ.get("B2_FUSED_NCOMM", "16")))
    L.b2_fused_set_debug(None)


def trace_rs(M, K, Nn, tag):
    a = torch.randn(M, K, device=dev, dtype=torch.bfloat16); W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
    L.b2_fused_set_debug(None)
    for i in range(3):
        F.gemm_rs(a, W)
    torch.cuda.synchronize(); dist.barrier()
    L.b2_fused_set_debug(N.c_p(dbg.data_ptr()))
    dbg.zero_(); F.gemm_rs(a, W); torch.cuda.synchronize(); dist.barrier()
    show(f"gemm_rs {tag} M{M} K{K} N{Nn}")
    L.b2_fused_set_debug(None)


trace_ag(6144, 768, 1536, "fc1")
trace_ag(12288, 768, 1536, "fc1")
trace_rs(12288, 384, 768, "proj")
trace_rs(24576, 384, 768, "proj")
trace_rs(24576, 1536, 768, "fc2")
dist.barrier(); dist.destroy_process_group()
