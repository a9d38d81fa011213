"""Host-side logic that needs no GPU: fake (meta) kernels, the linear-method mirrors' shape and
layout handling, the data-parallel row sharding, and the tensor-parallel column / row composition
run for real at world_size 2 over gloo (an oracle-backed stand-in replaces the CUDA ops there --
tests are allowed to call the oracle; the product never does)."""
from __future__ import annotations

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, bf16_bits, mk_int8_problem

import sglang_diffusion_b200 as b2
from oracle import restate


# ---------------------------------------------------------------- fake kernels ----

def test_fake_kernels_give_reference_shapes():
    M, N, K = 7, 48, 64
    a = torch.empty(M, K, dtype=torch.int8, device="meta")
    w = torch.empty(N, K, dtype=torch.int8, device="meta").t()
    sa = torch.empty(M, 1, device="meta")
    sb = torch.empty(N, 1, device="meta")
    out = torch.ops.sgl_kernel.int8_scaled_mm(a, w, sa, sb, torch.bfloat16, None)
    assert out.shape == (M, N) and out.dtype == torch.bfloat16
    f8 = torch.empty(M, K, dtype=torch.float8_e4m3fn, device="meta")
    w8 = torch.empty(N, K, dtype=torch.float8_e4m3fn, device="meta").t()
    out = torch.ops.sgl_kernel.fp8_scaled_mm(f8, w8, sa, sb, torch.float16, None)
    assert out.shape == (M, N) and out.dtype == torch.float16
    out = torch.ops.sgl_kernel.fp8_blockwise_scaled_mm(f8, w8, sa, sb, torch.bfloat16)
    assert out.shape == (M, N)
    x = torch.empty(M, K, dtype=torch.bfloat16, device="meta")
    q, s = torch.ops.sgl_kernel.per_token_quant_int8_cpu(x)
    assert q.shape == (M, K) and q.dtype == torch.int8 and s.shape == (M,) and s.dtype == torch.float32
    wn = torch.empty(N, K, dtype=torch.int8, device="meta")
    out = torch.ops.sgl_kernel.int8_scaled_mm_with_quant(x, wn, sb, None, torch.bfloat16, False)
    assert out.shape == (M, N) and out.dtype == torch.bfloat16


def test_group_scale_allocation_layouts():
    """fp8_kernel.py:435-476: column-major scales are logical [M, K/128] with stride(0) == 1."""
    s = b2.ops.create_per_token_group_quant_fp8_output_scale((10, 512), "cpu", 128, column_major_scales=True)
    assert s.shape == (10, 4) and s.stride() == (1, 10)
    s = b2.ops.create_per_token_group_quant_fp8_output_scale((10, 512), "cpu", 128, column_major_scales=True,
                                                            scale_tma_aligned=True)
    assert s.shape == (10, 4) and s.stride() == (1, 12)  # M padded to a multiple of 4
    s = b2.ops.create_per_token_group_quant_fp8_output_scale((10, 512), "cpu", 128, column_major_scales=False)
    assert s.shape == (10, 4) and s.is_contiguous()


def test_out_of_scope_modes_raise():
    """What is still not built says so: UE8M0 group scales.  (cal_sum, fused silu_and_mul and the
    static input_scale branch are built in round 2 -- tests/test_gpu_fused.py.)"""
    with pytest.raises(NotImplementedError):
        b2.ops.create_per_token_group_quant_fp8_output_scale((4, 256), "cpu", 128, True, True, scale_ue8m0=True)


def test_triton_quantizer_formula_delta_is_ties_and_zero_rows_only():
    """ADVICE: ops.per_token_quant_int8 carries the Triton quantizer's name (int8_kernel.py:28-89) but computes the
    CPU path's arithmetic (eps 1e-7, round-half-even; BASELINE makes that the bit-exact target).  Bound the delta
    against the Triton formula (eps 1e-10, round-half-away): codes differ only on exact .5 ties, by one LSB; scales
    differ only on rows whose absmax is below 1e-7."""
    from oracle import restate

    g = torch.Generator().manual_seed(0)
    x = (torch.randn(64, 512, generator=g) / 10).bfloat16()
    x[3] = 0  # all-zero row
    x[5, :8] = torch.tensor([0.5, 1.5, 2.5, -0.5, -1.5, 127.0, -127.0, 63.5]).bfloat16()  # ties once amax = 127
    x[5, 8:] = 0
    q_u8, s = restate.per_token_quant_int8_cpu(x)
    ours = q_u8.to(torch.int16) - 128
    xf = x.float()
    absmax = xf.abs().amax(-1, keepdim=True).clamp_min(1e-10)
    t = xf * (torch.tensor(127.0) / absmax)
    triton_q = torch.where(t >= 0, torch.floor(t + 0.5), torch.ceil(t - 0.5)).to(torch.int16)
    diff = (ours - triton_q).abs()
    assert int(diff.max()) <= 1
    ties = (t - torch.trunc(t)).abs() == 0.5
    assert bool((diff[~ties] == 0).all())
    assert int(diff[5].sum()) >= 1  # the tie row really exercises the difference
    triton_s = (absmax / 127).view(-1)
    differ = s != triton_s
    assert bool((differ == (xf.abs().amax(-1) < 1e-7)).all())


# -------------------------------------------------------------- linear methods ----

class _Layer(torch.nn.Module):
    pass


def test_int8_linear_method_weight_layout():
    m = b2.W8A8Int8LinearMethod()
    layer = _Layer()
    m.create_weights(layer, 64, [32, 16], 64, 48, torch.bfloat16)
    assert layer.weight.shape == (48, 64) and layer.weight.dtype == torch.int8
    assert layer.weight_scale.shape == (48, 1) and layer.weight_scale.dtype == torch.float32
    base = layer.weight.data_ptr()
    m.process_weights_after_loading(layer)
    # w8a8_int8.py:160-168: a transposed VIEW -- [K,N] with stride(0)==1, same memory
    assert layer.weight.shape == (64, 48) and layer.weight.stride() == (1, 64) and layer.weight.data_ptr() == base


def test_fp8_methods_weight_layout():
    m = b2.W8A8Fp8LinearMethod()
    layer = _Layer()
    m.create_weights(layer, 128, [256], 128, 256, torch.bfloat16)
    m.process_weights_after_loading(layer)
    assert layer.weight.dtype == torch.float8_e4m3fn and layer.weight.shape == (128, 256) and layer.weight.stride(0) == 1
    mb = b2.W8A8BlockFp8LinearMethod()
    layer = _Layer()
    mb.create_weights(layer, 512, [384], 512, 384, torch.bfloat16)
    assert layer.weight.shape == (384, 512) and layer.weight_scale_inv.shape == (3, 4)


# ---------------------------------------------------------------- row sharding ----

@pytest.mark.parametrize("M,world", [(32768, 8), (32768, 1), (10, 4), (3, 8), (0, 2), (8191, 2)])
def test_shard_rows_partitions_exactly(M, world):
    blocks = [b2.shard_rows(M, r, world) for r in range(world)]
    assert blocks[0][0] == 0 and blocks[-1][1] == M
    for (a0, a1), (b0, b1) in zip(blocks, blocks[1:]):
        assert a1 == b0
    sizes = [b - a for a, b in blocks]
    assert max(sizes) - min(sizes) <= 1 and sum(sizes) == M


# ------------------------------------------------- world_size-2 gloo composition ----

class OracleInt8Method(b2.W8A8Int8LinearMethod):
    """Test stand-in: same create_weights / process_weights_after_loading as the product method, but
    apply() computes with the CPU oracle so the TP composition can run without a GPU."""

    def apply(self, layer, x, bias=None):
        W = layer.weight.t().contiguous()  # back to [N,K]
        out = restate.int8_scaled_mm_with_quant(x.view(-1, x.shape[-1]).contiguous(), W, layer.weight_scale.view(-1),
                                                None if bias is None else bias.float())
        return out.view(*x.shape[:-1], W.shape[0])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _tp_worker(rank, world, port, outdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        M, N, K = 6, 64, 256
        x, W, Bs, bias = mk_int8_problem(M, N, K, seed=77)
        # -- row-parallel: W split on K, x already K-sharded, ONE all-reduce, bias on rank 0 only
        row = b2.QuantLinear(K, N, OracleInt8Method(), bias=True, parallel="row", tp_rank=rank, tp_size=world)
        row.weight.data.copy_(row.shard_of(W))
        row.weight_scale.data.copy_(Bs.view(-1, 1))
        row.bias.data.copy_(bias)
        row.quant_method.process_weights_after_loading(row)
        k0 = rank * (K // world)
        out_row = row(x[:, k0:k0 + K // world].contiguous())
        # -- column-parallel: W split on N, x replicated, no collective
        col = b2.QuantLinear(K, N, OracleInt8Method(), bias=True, parallel="column", tp_rank=rank, tp_size=world)
        col.weight.data.copy_(col.shard_of(W))
        col.weight_scale.data.copy_(col.shard_of(Bs.view(-1, 1)))
        col.bias.data.copy_(col.shard_of(bias))
        col.quant_method.process_weights_after_loading(col)
        out_col = col(x)
        # -- data-parallel rows: weights replicated, no collective
        r0, r1 = b2.shard_rows(M, rank, world)
        rep = b2.QuantLinear(K, N, OracleInt8Method(), bias=True)
        rep.weight.data.copy_(W)
        rep.weight_scale.data.copy_(Bs.view(-1, 1))
        rep.bias.data.copy_(bias)
        rep.quant_method.process_weights_after_loading(rep)
        out_dp = rep(x[r0:r1])
        torch.save({"row": out_row, "col": out_col, "dp": out_dp, "rows": (r0, r1)}, os.path.join(outdir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_tp2_and_dp2_composition_over_gloo(tmp_path):
    world = 2
    port = _free_port()
    mp.spawn(_tp_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    M, N, K = 6, 64, 256
    x, W, Bs, bias = mk_int8_problem(M, N, K, seed=77)
    full = restate.int8_scaled_mm_with_quant(x, W, Bs, bias.float())

    # column shards concatenate to the unsharded result bit-for-bit (same row scales on every rank)
    cat = torch.cat([r["col"] for r in res], dim=1)
    assert torch.equal(bf16_bits(cat), bf16_bits(full))
    # data-parallel row blocks concatenate to the unsharded result bit-for-bit
    dp = torch.cat([r["dp"] for r in res], dim=0)
    assert torch.equal(bf16_bits(dp), bf16_bits(full))
    assert [r["rows"] for r in res] == [(0, 3), (3, 6)]
    # row-parallel: every rank holds the same all-reduced tensor = bf16 sum of the per-shard outputs
    # (per-token scales are local to each K shard, bias only on rank 0 -- linear.py:1480-1490)
    assert torch.equal(bf16_bits(res[0]["row"]), bf16_bits(res[1]["row"]))
    parts = []
    for r in range(world):
        k0 = r * (K // world)
        parts.append(restate.int8_scaled_mm_with_quant(x[:, k0:k0 + K // world].contiguous(),
                                                       W[:, k0:k0 + K // world].contiguous(), Bs,
                                                       bias.float() if r == 0 else None))
    want = parts[0].clone()
    want += parts[1]
    assert torch.equal(bf16_bits(res[0]["row"]), bf16_bits(want))
    # and it approximates the unsharded layer (different quantization grid per shard)
    torch.testing.assert_close(res[0]["row"].float(), full.float(), rtol=5e-2, atol=5e-2)


# ------------------------------ row-parallel layer + norm + next quantizer, fallback composition over gloo ----

def _oracle_rmsnorm_quant(x, weight, eps=1e-6, residual=None, mode="int8", return_normed=False):
    """CPU stand-in for ops.rmsnorm_quant (oracle restatements): residual updated in place, int8 rows."""
    assert mode == "int8"
    y, new_res = restate.rmsnorm(x, weight, eps, residual)
    if residual is not None:
        residual.copy_(new_res)
    q_u8, s = restate.per_token_quant_int8_cpu(y.contiguous())
    q = (q_u8.to(torch.int16) - 128).to(torch.int8)
    return (q, s.view(-1, 1), y) if return_normed else (q, s.view(-1, 1))


def _norm_chain_worker(rank, world, port, outdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sglang_diffusion_b200 import linear as lin

        lin.ops.rmsnorm_quant = _oracle_rmsnorm_quant  # this process only
        M, N, K = 6, 64, 256
        x, W, Bs, bias = mk_int8_problem(M, N, K, seed=78)
        g = torch.Generator().manual_seed(5)
        wn = (torch.rand(N, generator=g) + 0.5).bfloat16()
        res_full = torch.randn(M, N, generator=g).bfloat16()
        row = b2.QuantLinear(K, N, OracleInt8Method(), bias=True, parallel="row", tp_rank=rank, tp_size=world)
        row.weight.data.copy_(row.shard_of(W))
        row.weight_scale.data.copy_(Bs.view(-1, 1))
        row.bias.data.copy_(bias)
        row.quant_method.process_weights_after_loading(row)
        k0 = rank * (K // world)
        xs = x[:, k0:k0 + K // world].contiguous()
        rows = M // world
        summed = row(xs).clone()
        rs = row.forward_reduce_scatter(xs)
        r_ar = res_full.clone()
        q_ar, s_ar = row.forward_allreduce_rmsnorm_quant(xs, wn, 1e-6, residual=r_ar, mode="int8")
        r_sp = res_full[rank * rows:(rank + 1) * rows].clone()
        q_sp, s_sp = row.forward_reduce_scatter_rmsnorm_quant(xs, wn, 1e-6, residual_local=r_sp, mode="int8")
        torch.save({"summed": summed, "rs": rs, "q_ar": q_ar, "s_ar": s_ar, "r_ar": r_ar, "q_sp": q_sp, "s_sp": s_sp,
                    "r_sp": r_sp}, os.path.join(outdir, f"n{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_row_parallel_norm_chain_fallbacks_over_gloo(tmp_path):
    """QuantLinear.forward_reduce_scatter / forward_allreduce_rmsnorm_quant / forward_reduce_scatter_rmsnorm_quant
    without a peer-memory communicator (the composition every rank count and dtype can fall back to): reduce-scatter
    rows equal the all-reduce's rows; the all-reduce form equals norm + quantizer of the summed rows on every rank; the
    sequence-parallel form gathers exactly the same codes and scales, and each rank's residual rows match."""
    world = 2
    mp.spawn(_norm_chain_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"n{r}.pt")) for r in range(world)]
    M, N = 6, 64
    rows = M // world
    g = torch.Generator().manual_seed(5)
    wn = (torch.rand(N, generator=g) + 0.5).bfloat16()
    res_full = torch.randn(M, N, generator=g).bfloat16()
    summed = res[0]["summed"]
    assert torch.equal(bf16_bits(summed), bf16_bits(res[1]["summed"]))
    want_res = res_full.clone()
    want_q, want_s = _oracle_rmsnorm_quant(summed, wn, 1e-6, residual=want_res)
    for r in range(world):
        assert torch.equal(bf16_bits(res[r]["rs"]), bf16_bits(summed[r * rows:(r + 1) * rows]))
        assert torch.equal(res[r]["q_ar"], want_q) and torch.equal(res[r]["s_ar"], want_s)
        assert torch.equal(bf16_bits(res[r]["r_ar"]), bf16_bits(want_res))
        assert torch.equal(res[r]["q_sp"], want_q) and torch.equal(res[r]["s_sp"], want_s)
        assert torch.equal(bf16_bits(res[r]["r_sp"]), bf16_bits(want_res[r * rows:(r + 1) * rows]))
