"""-m gpu parity tests of the W8A8-int8 MoE path: block alignment, grouped GEMM bit-exact against
the dense int8 op per expert, fused_experts against the restatement of moe_int8.cpp."""
from __future__ import annotations

import pytest
import torch

import sglang_diffusion_b200 as b2
from sglang_diffusion_b200 import capi, moe, ops
from oracle import restate  # checker only

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _routing(T, E, topk, seed, skew=False):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(T, E, generator=g)
    if skew:
        logits[:, 0] += 4.0  # one hot expert, several cold (possibly empty) ones
        logits[:, E - 1] -= 50.0
    w, ids = torch.topk(torch.softmax(logits, -1), topk, dim=-1)
    return (w / w.sum(-1, keepdim=True)).float(), ids.to(torch.int32)


@pytest.mark.parametrize("T,E,topk", [(1, 8, 2), (7, 8, 2), (300, 16, 4), (1024, 64, 8), (33, 256, 8)])
@pytest.mark.parametrize("skew", [False, True])
def test_moe_align_block128_invariants(T, E, topk, skew):
    _, ids = _routing(T, E, topk, T + E, skew)
    sorted_ids, tile_expert, inv_pos, npp = moe.moe_align_block128(ids.to(DEV), E)
    sorted_ids, tile_expert, inv_pos, npp = sorted_ids.cpu(), tile_expert.cpu(), inv_pos.cpu(), int(npp.item())
    n = T * topk
    flat = ids.view(-1)
    counts = torch.bincount(flat.long(), minlength=E)
    assert npp == int(((counts + 127) // 128 * 128).sum()) and npp % 128 == 0
    assert bool((tile_expert[npp // 128:] == -1).all()) and bool((sorted_ids[npp:] == n).all())
    # every flat index appears exactly once, at inv_pos, inside a block of its own expert
    real = sorted_ids[sorted_ids < n]
    assert real.numel() == n and torch.equal(torch.sort(real).values, torch.arange(n, dtype=torch.int32))
    assert torch.equal(sorted_ids[inv_pos.long()], torch.arange(n, dtype=torch.int32))
    assert torch.equal(tile_expert[(inv_pos // 128).long()], flat)
    # blocks of one expert are contiguous and experts ascend
    used = tile_expert[: npp // 128]
    assert bool((used[1:] >= used[:-1]).all()) and bool((used >= 0).all())


@pytest.mark.parametrize("T,E,topk,H,N", [(5, 4, 2, 512, 256), (300, 8, 2, 1024, 768), (64, 16, 4, 2048, 1408),
                                          (700, 8, 2, 544, 384)])
def test_grouped_int8_gemm_equals_dense_per_expert_bitwise(T, E, topk, H, N):
    """Rows of one expert through the grouped kernel == the dense int8_scaled_mm with that expert's weight (int32
    accumulators are exact, the epilogue arithmetic is the same function): bit for bit, incl. the oracle."""
    _, ids = _routing(T, E, topk, T + H, skew=True)
    g = torch.Generator().manual_seed(H)
    x = (torch.randn(T, H, generator=g) / 10).bfloat16()
    w = torch.randint(-127, 128, (E, N, H), dtype=torch.int8, generator=g)
    ws = torch.rand(E, N, generator=g) * 1e-2
    sorted_ids, tile_expert, inv_pos, npp = moe.moe_align_block128(ids.to(DEV), E)
    P = sorted_ids.numel()
    a_q = torch.empty(P, H, dtype=torch.int8, device=DEV)
    a_s = torch.empty(P, dtype=torch.float32, device=DEV)
    xd = x.to(DEV)
    capi.check(capi.lib.b2_quant_int8_rows_gather(xd.data_ptr(), capi.DTYPE_BF16, H, sorted_ids.data_ptr(), topk,
                                                  T * topk, a_q.data_ptr(), H, a_s.data_ptr(), P, H,
                                                  torch.cuda.current_stream().cuda_stream))
    out = moe.grouped_int8_mm(a_q, a_s, w.to(DEV), ws.to(DEV), tile_expert, torch.bfloat16).cpu()
    inv = inv_pos.cpu().long().view(T, topk)
    # gathered quantization == quantization of the source row
    q_ref, s_ref = restate.per_token_quant_int8_cpu(x)
    assert torch.equal((a_q.cpu()[inv[:, 0]].to(torch.int16) + 128).to(torch.uint8), q_ref)
    assert torch.equal(a_s.cpu()[inv[:, 0]], s_ref)
    for e in range(E):
        tok, kk = (ids.long() == e).nonzero(as_tuple=True)
        if tok.numel() == 0:
            continue
        ref = restate.int8_scaled_mm_with_quant(x[tok].contiguous(), w[e].contiguous(), ws[e], None)
        got = out[inv[tok, kk]]
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16)), e
        dense = b2.int8_scaled_mm_with_quant(xd[tok.to(DEV)].contiguous(), w[e].to(DEV), ws[e].to(DEV), None,
                                             torch.bfloat16, False)
        assert torch.equal(got.view(torch.int16), dense.cpu().view(torch.int16))
    # padding rows inside the used blocks are exact zeros (all-zero codes); blocks past num_post_pad are skipped by
    # the kernel (tile_expert == -1) and their rows are never read
    used = int(npp.item())
    pad = sorted_ids.cpu()[:used] >= T * topk
    assert bool((out[:used][pad] == 0).all())


@pytest.mark.parametrize("T,E,topk,H,I", [(1, 8, 2, 512, 256), (37, 8, 2, 1024, 512), (256, 16, 4, 2048, 768),
                                          (1000, 32, 8, 1024, 384)])
def test_fused_experts_int8_vs_restatement(T, E, topk, H, I):
    """End to end against the restatement of moe_int8.cpp.  The GEMM / quantizer stages are bit-exact (test above);
    silu uses the sm_100 tanh.approx, so activations can differ by 1 bf16 ulp before the second quantizer: the
    reference's own tolerance for this op (test/srt/cpu/test_moe.py, atol = rtol = 1e-2 on O(1) outputs) is applied
    relative to the output scale."""
    tw, ids = _routing(T, E, topk, T + E + H)
    g = torch.Generator().manual_seed(I)
    x = (torch.randn(T, H, generator=g) / 10).bfloat16()
    w13 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8, generator=g)
    w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8, generator=g)
    s13 = torch.rand(E, 2 * I, 1, generator=g) * 2e-3
    s2 = torch.rand(E, H, 1, generator=g) * 2e-3
    capi.reset_launch_count()
    out = b2.fused_experts(x.to(DEV), w13.to(DEV), w2.to(DEV), tw.to(DEV), ids.to(DEV), s13.to(DEV), s2.to(DEV))
    assert capi.launch_count() == 6  # align, gather+quant, GEMM, silu+quant, GEMM, combine
    ref = restate.fused_experts_int8(x, w13, w2, tw, ids, s13, s2)
    o, r = out.float().cpu(), ref.float()
    assert torch.isfinite(o).all()
    scale = r.abs().max().item()
    assert (o - r).abs().max().item() <= 1e-2 * scale + 1e-2 * 0
    assert ((o - r).norm() / r.norm()).item() < 5e-3
    # deterministic: a second call gives the same bits (atomics only order rows inside an expert's block range)
    out2 = b2.fused_experts(x.to(DEV), w13.to(DEV), w2.to(DEV), tw.to(DEV), ids.to(DEV), s13.to(DEV), s2.to(DEV))
    assert torch.equal(out.view(torch.int16), out2.view(torch.int16))


def test_moe_method_mirror_and_graph_capture():
    T, E, topk, H, I = 64, 8, 2, 512, 256
    tw, ids = _routing(T, E, topk, 3)
    m = b2.W8A8Int8MoEMethod()
    layer = torch.nn.Module()
    m.create_weights(layer, E, H, I, torch.bfloat16, device=DEV)
    assert layer.w13_weight.shape == (E, 2 * I, H) and layer.w2_weight.shape == (E, H, I)
    assert layer.w13_weight_scale.shape == (E, 2 * I, 1) and layer.w2_weight_scale.shape == (E, H, 1)
    g = torch.Generator(device=DEV).manual_seed(1)
    layer.w13_weight.data.copy_(torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8, generator=g, device=DEV))
    layer.w2_weight.data.copy_(torch.randint(-127, 128, (E, H, I), dtype=torch.int8, generator=g, device=DEV))
    layer.w13_weight_scale.data.mul_(1e-3)
    layer.w2_weight_scale.data.mul_(1e-3)
    m.process_weights_after_loading(layer)
    x = (torch.randn(T, H, generator=g, device=DEV) / 10).bfloat16()
    twd, idd = tw.to(DEV), ids.to(DEV)
    eager = m.apply(layer, x, twd, idd)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out = m.apply(layer, x, twd, idd)
    torch.cuda.current_stream().wait_stream(side)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int16), eager.view(torch.int16))


def test_fused_experts_f16_activations():
    """f16 hidden states (the other activation dtype of the linear methods): same routing and weights as a bf16 run,
    outputs agree to the activation rounding."""
    T, E, topk, H, I = 200, 8, 2, 1024, 512
    tw, ids = _routing(T, E, topk, 11)
    g = torch.Generator().manual_seed(4)
    x = (torch.randn(T, H, generator=g) / 10)
    w13 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8, generator=g).to(DEV)
    w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8, generator=g).to(DEV)
    s13 = (torch.rand(E, 2 * I, 1, generator=g) * 2e-3).to(DEV)
    s2 = (torch.rand(E, H, 1, generator=g) * 2e-3).to(DEV)
    o16 = b2.fused_experts(x.half().to(DEV), w13, w2, tw.to(DEV), ids.to(DEV), s13, s2)
    obf = b2.fused_experts(x.bfloat16().to(DEV), w13, w2, tw.to(DEV), ids.to(DEV), s13, s2)
    assert o16.dtype == torch.float16 and bool(torch.isfinite(o16).all())
    assert ((o16.float() - obf.float()).norm() / obf.float().norm()).item() < 2e-2
