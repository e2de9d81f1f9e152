"""Kernel oracle tests at the MODEL's shapes with ELEMENT-WISE bounds (round-1 review: the small-shape tests compared
max|err| against the tensor's max, which leaves small elements unchecked, and never wrapped the GEMM's smem ring
more than once).

GEMM error model, per output element: inputs are exact bf16, products are exact in fp32, the accumulation is fp32
(error ~ K * 2^-24 * sum|a||b|) and the result is rounded once to bf16 (2^-9 relative).  So

    |c - ref| <= 2^-8 * |ref| + 2e-6 * K^0.5 * (|A| @ |B|)          (first term: rounding, second: accumulation order)

is checked for EVERY element, with |A| @ |B| computed in fp32 as well.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from llm_training_b200.ops import reference as ref
from llm_training_b200.ops.adamw_op import adamw_step
from llm_training_b200.ops.attention_op import packed_attention
from llm_training_b200.ops.gemm_op import gemm

DEV = 'cuda'


@pytest.fixture(autouse=True)
def _exact_fp32_matmul():
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False       # the oracle must be real fp32
    yield
    torch.backends.cuda.matmul.allow_tf32 = prev


def _assert_elementwise(c: torch.Tensor, want: torch.Tensor, abs_prod: torch.Tensor, k: int, out_bf16: bool = True):
    err = (c.float() - want).abs()
    bound = (2.0 ** -8 if out_bf16 else 2.0 ** -20) * want.abs() + 2e-6 * (k ** 0.5) * abs_prod + 1e-30
    worst = (err / bound).max().item()
    assert worst <= 1.0, f'element-wise bound exceeded by {worst:.2f}x at {int((err / bound).argmax())}'


# (M, N, K): q|k|v, gate|up, down, o_proj of LLaMA-3-8B at 8192 tokens; M / N tails on top of a deep K
MODEL_SHAPES = [(8192, 6144, 4096), (8192, 28672, 4096), (8192, 4096, 14336), (8192, 4096, 4096),
                (8192 - 72, 6144 + 40, 4096)]


@pytest.mark.parametrize('cg', [1, 2])
@pytest.mark.parametrize('amn,bmn', [(0, 0), (0, 1), (1, 1)])       # forward, dgrad, wgrad operand layouts
@pytest.mark.parametrize('m,n,k', MODEL_SHAPES)
def test_gemm_model_shapes_elementwise(cg, amn, bmn, m, n, k):
    g = torch.Generator(device=DEV).manual_seed(m * 7 + n * 3 + k)
    a = (torch.randn((k, m) if amn else (m, k), device=DEV, generator=g) * 0.5).bfloat16()
    b = (torch.randn((k, n) if bmn else (n, k), device=DEV, generator=g) * 0.5).bfloat16()
    A = a.float().t() if amn else a.float()
    B = b.float() if bmn else b.float().t()
    c = gemm(a, b, a_mn_major=bool(amn), b_mn_major=bool(bmn), cta_group=cg)
    _assert_elementwise(c, A @ B, A.abs() @ B.abs(), k)


@pytest.mark.parametrize('cg', [1, 2])
def test_gemm_wgrad_shape_fp32_accumulate(cg):
    """dW[4096, 14336] += X^T dY over 8192 tokens, accumulated in place into an fp32 buffer (direct wgrad)."""
    m, n, k = 4096, 14336, 8192
    g = torch.Generator(device=DEV).manual_seed(5)
    a = (torch.randn(k, m, device=DEV, generator=g) * 0.5).bfloat16()       # [K, M]: A is MN-major
    b = (torch.randn(k, n, device=DEV, generator=g) * 0.5).bfloat16()       # [K, N]: B is MN-major
    acc = torch.randn(m, n, device=DEV, generator=g)
    want = acc + a.float().t() @ b.float()
    prod = a.float().t().abs() @ b.float().abs() + acc.abs()
    gemm(a, b, a_mn_major=True, b_mn_major=True, out=acc, residual=acc, cta_group=cg)
    _assert_elementwise(acc, want, prod, k, out_bf16=False)


def test_gemm_lm_head_shape():
    """Logits chunk of the fused linear + loss: N = 128256 (not a multiple of 256), K = 4096."""
    m, n, k = 2048, 128256, 4096
    g = torch.Generator(device=DEV).manual_seed(6)
    a = (torch.randn(m, k, device=DEV, generator=g) * 0.5).bfloat16()
    b = (torch.randn(n, k, device=DEV, generator=g) * 0.05).bfloat16()
    c = gemm(a, b, cta_group=2)
    _assert_elementwise(c, a.float() @ b.float().t(), a.float().abs() @ b.float().abs().t(), k)


def _attention_oracle(qkv, start, hq, hkv, d, window, dout):
    """fp32 oracle on the GPU, one kv group at a time (the [Hq, T, T] score tensor of all heads would not fit)."""
    tokens = qkv.shape[0]
    ref_in = qkv.detach().float().requires_grad_()
    q, k, v = ref_in.split([hq * d, hkv * d, hkv * d], dim=-1)
    q, k, v = q.reshape(tokens, hq, d), k.reshape(tokens, hkv, d), v.reshape(tokens, hkv, d)
    group = hq // hkv
    outs = []
    for kv_head in range(hkv):
        o = ref.attention_ref(q[:, kv_head * group:(kv_head + 1) * group], k[:, kv_head:kv_head + 1],
                              v[:, kv_head:kv_head + 1], start, None, window)
        o.backward(dout.float().view(tokens, hq, d)[:, kv_head * group:(kv_head + 1) * group], retain_graph=False)
        outs.append(o.detach())
    return torch.cat(outs, dim=1).reshape(tokens, hq * d), ref_in.grad


@pytest.mark.parametrize('tokens,hq,hkv,doc_lens,window', [
    (8192, 8, 2, [3000, 1192, 2500, 1500], None),        # >= 3 documents in an 8192-token pack
    (8192, 4, 1, [5000, 3192], 1024),                    # sliding window + documents
    (8192, 4, 2, None, None),                            # one dense causal sequence (the pre-training shape)
    (32768, 2, 1, [20000, 700, 12068], 4096),            # long pack
])
def test_attention_long_sequences_vs_fp32_oracle(tokens, hq, hkv, doc_lens, window):
    d = 128
    g = torch.Generator(device=DEV).manual_seed(tokens + hq)
    qkv = (torch.randn(tokens, (hq + 2 * hkv) * d, device=DEV, generator=g) * 0.5).bfloat16().requires_grad_()
    if doc_lens is None:
        start = torch.zeros(tokens, dtype=torch.int32, device=DEV)
        end = torch.full((tokens,), tokens, dtype=torch.int32, device=DEV)
    else:
        assert sum(doc_lens) == tokens
        s, e, pos = [], [], 0
        for n in doc_lens:
            s += [pos] * n
            e += [pos + n] * n
            pos += n
        start = torch.tensor(s, dtype=torch.int32, device=DEV)
        end = torch.tensor(e, dtype=torch.int32, device=DEV)
    dout = torch.randn(tokens, hq * d, device=DEV, generator=g).bfloat16()
    out = packed_attention(qkv, start, end, hq, hkv, d, None, window)
    out.backward(dout)
    o_ref, g_ref = _attention_oracle(qkv, start, hq, hkv, d, window, dout)
    # outputs are convex combinations of V rows (|v| <~ 2.5): bf16 rounding of P (2^-8) and of O
    err = (out.float() - o_ref).abs()
    assert (err <= 2.0 ** -6 * o_ref.abs() + 6e-3).all(), f'out: {err.max().item()}'
    width = qkv.shape[1]
    grads = qkv.grad.float()
    for name, sl in (('dq', slice(0, hq * d)), ('dk', slice(hq * d, (hq + hkv) * d)),
                     ('dv', slice((hq + hkv) * d, width))):
        e_abs = (grads[:, sl] - g_ref[:, sl]).abs()
        # per-row scale: a gradient row is a sum over the visible keys / queries; compare against the row's own norm
        row_scale = g_ref[:, sl].reshape(tokens, -1, d).norm(dim=-1, keepdim=True).expand(-1, -1, d).reshape(tokens, -1)
        bound = 2.0 ** -6 * g_ref[:, sl].abs() + 1.5e-2 * row_scale / d ** 0.5 + 2e-3
        worst = (e_abs / bound).max().item()
        assert worst <= 1.0, f'{name}: bound exceeded by {worst:.2f}x (max abs err {e_abs.max().item():.4g})'


def test_adamw_single_step_against_float64():
    """One update of the fused AdamW kernel against a float64 restatement with IDENTICAL inputs (fixed gradient and
    norm), so nothing compounds: masters must agree to a few fp32 ulps of the update's operands, the bf16 copies
    may differ only where the fp32 value sits within that distance of a rounding boundary."""
    n = 1 << 20
    g = torch.Generator(device=DEV).manual_seed(3)
    for step in (1, 2, 1000):
        master = torch.randn(n, device=DEV, generator=g)
        m = torch.randn(n, device=DEV, generator=g) * 1e-3
        v = torch.rand(n, device=DEV, generator=g) * 1e-5
        grad = (torch.randn(n, device=DEV, generator=g) * 1e-2).bfloat16()
        norm_sq = (grad.float() ** 2).sum().reshape(1)
        # the C ABI takes the hyper-parameters as fp32: the oracle uses the same (rounded) values in float64
        lr, b1, b2, eps, wd = (float(torch.tensor(x, dtype=torch.float32)) for x in (1e-3, 0.9, 0.999, 1e-8, 0.1))
        max_norm = 1.0
        p64, m64, v64 = master.double(), m.double(), v.double()
        clip = min(1.0, max_norm / (float(norm_sq.sqrt()) + 1e-6))
        g64 = grad.double() * clip
        m64 = b1 * m64 + (1 - b1) * g64
        v64 = b2 * v64 + (1 - b2) * g64 * g64
        delta = lr / (1 - b1 ** step) * m64 / (v64.sqrt() / (1 - b2 ** step) ** 0.5 + eps)
        p64 = p64 * (1 - lr * wd) - delta
        param = torch.empty(n, dtype=torch.bfloat16, device=DEV)
        adamw_step(master, m, v, grad, param, lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd, step=step,
                   max_norm=max_norm, norm_sq=norm_sq)
        # fp32 evaluation: a couple of roundings on p (2^-23 each) plus ~10 roundings inside the update term
        tol = 2.0 ** -22 * p64.abs() + 2e-6 * delta.abs() + 1e-12
        worst = ((master.double() - p64).abs() / tol).max().item()
        assert worst <= 1.0, f'step {step}: master differs from the float64 result by {worst:.2f}x the fp32 bound'
        # moments: (1 - beta) is itself rounded to fp32, then two products and a sum -> 4 roundings of 2^-24 each
        assert ((m.double() - m64).abs() <= 2.0 ** -21 * (m64.abs() + (1 - b1) * g64.abs()) + 1e-12).all()
        assert ((v.double() - v64).abs() <= 2.0 ** -21 * v64.abs() + 1e-16).all()
        flips = (param.float() != p64.float().bfloat16().float()).float().mean().item()
        assert flips < 1e-3, f'step {step}: {flips:.2e} of the bf16 copies differ from round(float64 result)'
