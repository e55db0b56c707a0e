"""Two ranks over NCCL (needs >= 2 GPUs on the box; skipped otherwise): the vocabulary-sharded
lens and LL-Top-k of dist.py give the single-GPU readout of the same rows."""
import os
import socket

import pytest
import torch

from _common import cuda_bf16, synthetic

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as td
    from taboo_brittleness_b200 import dist as tbd
    from taboo_brittleness_b200 import ops
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    td.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        P, T, d, V = 4, 128, 512, 8192
        x = synthetic.make_residuals(P, T, d, seed=21)
        lens = synthetic.make_lens(d, V, seed=21)
        sid = synthetic.make_secret_ids(P, T, V, seed=21).cuda()
        synthetic.plant_secret_direction(lens, x, sid.cpu())
        w_norm, wu = lens.w_norm.cuda(), cuda_bf16(lens.W_U)
        h = ops.rmsnorm(cuda_bf16(x), w_norm)
        cur = torch.randint(0, V, (P * T,), generator=torch.Generator().manual_seed(3), dtype=torch.int32).cuda()
        full = ops.lens_merge(ops.lens_partials(h, wu, softcap=30.0, secret_id=sid, cta_pair=1))
        full_v, full_i = ops.ll_topk(h, wu, full.lse, cur, None, softcap=30.0, cta_pair=1)
        v0, v1 = tbd.vocab_shard(V, rank, world)
        shard = ops.LensWeights(w_norm, wu[v0:v1].contiguous(), vocab_offset=v0)
        out = tbd.lens_readout_vocab_sharded(h, shard, softcap=30.0, secret_id=sid, cta_pair=4)
        sv, si = tbd.ll_topk_vocab_sharded(h, shard, out.lse, cur, None, softcap=30.0, cta_pair=4)
        torch.cuda.synchronize()
        ok = (torch.equal(out.top_id, full.top_id) and torch.allclose(out.lse, full.lse, atol=1e-5)
              and torch.allclose(out.secret_prob, full.secret_prob, atol=1e-6)
              and torch.equal(si, full_i) and torch.allclose(sv, full_v, rtol=1e-5, atol=1e-7))
        q.put((rank, bool(ok)))
    finally:
        td.destroy_process_group()


def test_vocab_sharded_lens_and_ll_topk_over_nccl(cuda_lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
    results = sorted(q.get(timeout=5) for _ in range(2))
    assert results == [(0, True), (1, True)]
