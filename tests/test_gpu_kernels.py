"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op.

Inputs of tensor-core kernels are rounded to bf16 first and the reference runs in fp32 on those rounded values,
so the only difference left is accumulation order (and bf16 rounding of bf16 outputs).
"""

import pytest
import torch

from tensorflow_distributed_b200.ops import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from tensorflow_distributed_b200.ops import native
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    native.lib()  # fail loudly if the .so is missing
    native.set_device(torch.cuda.current_device())
    return native


def bf(x):
    return x.to(torch.bfloat16)


def rel_err(a, b):
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


DEV = "cuda"


# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K,bn,splits", [(128, 1024, 3136, 128, 16), (128, 1024, 3136, 128, 1), (100, 72, 200, 64, 1),
                                             (256, 512, 128, 256, 1), (1000, 1024, 3136, 128, 2)])
def test_gemm_fwd_kmajor_a_mnmajor_b(nat, M, N, K, bn, splits):
    torch.manual_seed(0)
    A = bf(torch.randn(M, K, device=DEV))
    W = bf(torch.randn(K, N, device=DEV))
    ref = A.float() @ W.float()
    out = torch.zeros(M, N, device=DEV)
    mode = nat.EPI_ATOMIC_F32 if splits > 1 else nat.EPI_STORE_F32
    nat.gemm_bf16(A, 0, W, 1, M, N, K, out, mode=mode, bn=bn, splits=splits)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 2e-3
    if splits > 1:  # split-K without atomics: partial planes, summed by the consumer
        planes = torch.zeros(splits, M, N, device=DEV)
        nat.gemm_bf16(A, 0, W, 1, M, N, K, planes, mode=nat.EPI_STORE_F32, bn=bn, splits=splits, split_stride=M * N,
                      ld_out=N)
        torch.cuda.synchronize()
        assert rel_err(planes.sum(0), ref) < 2e-3
    # the CUDA-core reference kernel agrees too
    out2 = torch.zeros(M, N, device=DEV)
    nat.gemm_ref(A, W, M, N, K, out2)
    assert rel_err(out2, ref) < 2e-3


def test_gemm_dgrad_both_kmajor(nat):
    torch.manual_seed(1)
    B_, H, Kin = 128, 1024, 3136
    dy = bf(torch.randn(B_, H, device=DEV))
    W = bf(torch.randn(Kin, H, device=DEV))  # [in, out]; dgrad reads it as [N=in, K=out], K contiguous
    ref = dy.float() @ W.float().t()
    out = torch.zeros(B_, Kin, device=DEV)
    nat.gemm_bf16(dy, 0, W, 0, B_, Kin, H, out, mode=nat.EPI_ATOMIC_F32, bn=64, splits=3)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 2e-3


@pytest.mark.parametrize("Bsz", [128, 96])
def test_gemm_wgrad_both_mnmajor(nat, Bsz):
    torch.manual_seed(2)
    Kin, H = 3136, 1024
    x = bf(torch.randn(Bsz, Kin, device=DEV))
    dy = bf(torch.randn(Bsz, H, device=DEV))
    ref = x.float().t() @ dy.float()
    out = torch.zeros(Kin, H, device=DEV)
    nat.gemm_bf16(x, 1, dy, 1, Kin, H, Bsz, out, mode=nat.EPI_STORE_F32, bn=128)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 2e-3


def test_gemm_mnmajor_a_kmajor_b(nat):
    torch.manual_seed(3)
    M, N, K = 256, 192, 320
    At = bf(torch.randn(K, M, device=DEV))
    Bt = bf(torch.randn(N, K, device=DEV))
    ref = At.float().t() @ Bt.float().t()
    out = torch.zeros(M, N, device=DEV)
    nat.gemm_bf16(At, 1, Bt, 0, M, N, K, out, mode=nat.EPI_STORE_F32, bn=64)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 2e-3


def test_gemm_epilogues_bias_relu_and_relumask_and_padded_ld(nat):
    torch.manual_seed(4)
    M, N, K, ld = 128, 100, 784, 104
    A = bf(torch.randn(M, K, device=DEV))
    W = torch.zeros(K, ld, device=DEV, dtype=torch.bfloat16)
    W[:, :N] = bf(torch.randn(K, N, device=DEV))
    bias = torch.randn(N, device=DEV)
    ref = torch.relu(A.float() @ W[:, :N].float() + bias)
    out = torch.full((M, ld), 7.0, device=DEV, dtype=torch.bfloat16)
    nat.gemm_bf16(A, 0, W, 1, M, N, K, out, mode=nat.EPI_BIAS_RELU_BF16, bn=64, bias=bias)
    torch.cuda.synchronize()
    assert rel_err(out[:, :N], ref) < 1e-2
    assert torch.all(out[:, N:] == 7.0)  # pad columns untouched
    # dgrad through a ReLU: mask from an activation tensor
    dy = bf(torch.randn(M, N, device=DEV))
    dyp = torch.zeros(M, ld, device=DEV, dtype=torch.bfloat16)
    dyp[:, :N] = dy
    act = bf(torch.randn(M, K, device=DEV))
    ref2 = (dy.float() @ W[:, :N].float().t()) * (act.float() > 0)
    out2 = torch.zeros(M, K, device=DEV, dtype=torch.bfloat16)
    nat.gemm_bf16(dyp, 0, W, 0, M, K, N, out2, mode=nat.EPI_RELUMASK_BF16, bn=64, relu_mask=act)
    torch.cuda.synchronize()
    assert rel_err(out2, ref2) < 1e-2


# ------------------------------------------------------------------------------------------------
# tf32 arm: fp32 operands in memory, tcgen05 kind::tf32.  Reference = fp32 matmul of the tf32-truncated operands
# (the tensor core drops the low 13 mantissa bits), so what is left is accumulation order.
def tf32_trunc(x):
    return (x.contiguous().view(torch.int32) & ~0x1fff).view(torch.float32)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K,bn,splits", [(128, 1024, 3136, 64, 8), (100, 72, 200, 64, 1), (256, 512, 128, 256, 1),
                                             (3136, 1024, 128, 128, 1)])
def test_gemm_tf32_all_major_combinations(nat, a_mn, b_mn, M, N, K, bn, splits):
    if a_mn == 1 and M % 4:
        M = (M + 3) // 4 * 4  # leading dimension of an MN-major fp32 operand must keep 16-byte rows
    torch.manual_seed(7)
    A = torch.randn(M, K, device=DEV)
    Bm = torch.randn(K, N, device=DEV)
    ref = tf32_trunc(A) @ tf32_trunc(Bm)
    ref64 = (tf32_trunc(A).double() @ tf32_trunc(Bm).double()).float()
    As = A.t().contiguous() if a_mn else A                # a_mn: stored [K, M]
    Bs = Bm if b_mn else Bm.t().contiguous()              # b_mn=0: stored [N, K]
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        out = torch.full((max(1, splits), M, N), 7.0, device=DEV)
        nat.gemm_bf16(As, a_mn, Bs, b_mn, M, N, K, out, mode=nat.EPI_STORE_F32, bn=bn, splits=splits,
                      split_stride=M * N, ld_out=N, tf32=True)
        torch.cuda.synchronize()
        got = out.sum(0)
        assert rel_err(got, ref64) < 2e-5, (a_mn, b_mn)
        # and it is tf32, not bf16 and not fp32: against the untruncated fp32 product the error is ~1e-3, not ~1e-2
        full = (A.double() @ Bm.double()).float()
        assert rel_err(got, full) < 4e-3
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def test_gemm_tf32_epilogues_bias_relu_and_relumask_f32(nat):
    torch.manual_seed(8)
    M, N, K, ld = 128, 100, 784, 104
    A = torch.randn(M, K, device=DEV)
    W = torch.zeros(K, ld, device=DEV)
    W[:, :N] = torch.randn(K, N, device=DEV)
    bias = torch.randn(N, device=DEV)
    ref = torch.relu((tf32_trunc(A).double() @ tf32_trunc(W[:, :N]).double()).float() + bias)
    out = torch.full((M, ld), 7.0, device=DEV)
    nat.gemm_bf16(A, 0, W, 1, M, N, K, out, mode=nat.EPI_BIAS_RELU_F32, bn=64, bias=bias, tf32=True)
    torch.cuda.synchronize()
    assert rel_err(out[:, :N], ref) < 2e-5
    assert torch.all(out[:, N:] == 7.0)
    dy = torch.zeros(M, ld, device=DEV)
    dy[:, :N] = torch.randn(M, N, device=DEV)
    act = torch.randn(M, K, device=DEV)
    ref2 = (tf32_trunc(dy[:, :N]).double() @ tf32_trunc(W[:, :N]).double().t()).float() * (act > 0)
    out2 = torch.zeros(M, K, device=DEV)
    nat.gemm_bf16(dy, 0, W, 0, M, K, N, out2, mode=nat.EPI_RELUMASK_F32, bn=64, relu_mask=act, tf32=True)
    torch.cuda.synchronize()
    assert rel_err(out2, ref2) < 2e-5


# ------------------------------------------------------------------------------------------------
def _conv1_ref(x, w, b):
    y = torch.relu(oracle.conv2d_same_nhwc(x.reshape(-1, 28, 28, 1), w) + b)
    return oracle.maxpool2x2_nhwc(y), y


def _scratch(nat, kind, B):
    return torch.zeros(nat.scratch_floats(kind, B), device=DEV)


@pytest.mark.parametrize("tf32", [False, True])
@pytest.mark.parametrize("B", [128, 5])
def test_conv1_fwd_and_wgrad(nat, B, tf32):
    torch.manual_seed(5)
    adt = torch.float32 if tf32 else torch.bfloat16
    x = torch.rand(B, 784, device=DEV)
    w = torch.randn(5, 5, 1, 32, device=DEV)
    b = torch.randn(32, device=DEV)
    pool = torch.empty(B, 14, 14, 32, device=DEV, dtype=adt)
    am = torch.empty(B, 14, 14, 32, device=DEV, dtype=torch.uint8)
    nat.conv1_fwd(x, w, b, pool, am, B, tf32=tf32)
    torch.cuda.synchronize()
    ref, pre = _conv1_ref(x, w, b)
    assert rel_err(pool, ref) < (1e-5 if tf32 else 1e-2)
    assert int(am.max()) <= 3
    # backward: oracle gradient of sum(pool * dpool) w.r.t. w and b; dpool arrives as two partial planes
    dpool = torch.randn(B, 14, 14, 32, device=DEV)
    planes = torch.stack([0.25 * dpool, 0.75 * dpool])
    wl, bl = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    r2, _ = _conv1_ref(x, wl, bl)
    (r2 * dpool).sum().backward()
    scratch = _scratch(nat, "conv1_wgrad", B)
    res = []
    for _ in range(2):  # written, not accumulated; a second launch on the same scratch gives the same bits
        dw = torch.full((25, 32), 7.0, device=DEV)
        db = torch.full((32,), 7.0, device=DEV)
        nat.conv1_wgrad(x, planes, pool, am, dw, db, scratch, B, tf32=tf32, nsplit=2, split_stride=dpool.numel())
        torch.cuda.synchronize()
        res.append((dw, db))
    assert rel_err(res[0][0], wl.grad.reshape(25, 32)) < 2e-3
    assert rel_err(res[0][1], bl.grad) < 2e-3
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_conv1_fwd_gated_on_local_flags_equals_plain(nat):
    """The gated form (every block first waits until the shards' version words in LOCAL memory have reached the
    worker's step counter) computes exactly what the plain form computes; the flags are already satisfied here."""
    B = 37
    torch.manual_seed(8)
    x = torch.rand(B, 784, device=DEV)
    w = torch.randn(25, 32, device=DEV)
    b = torch.randn(32, device=DEV)
    outs = []
    for gated in (False, True):
        pool = torch.zeros(B, 14, 14, 32, device=DEV, dtype=torch.bfloat16)
        am = torch.zeros(B, 14, 14, 32, device=DEV, dtype=torch.uint8)
        if gated:
            flags = torch.tensor([7, 9, 8], dtype=torch.int32, device=DEV)  # versions of three shards
            want = torch.tensor([7], dtype=torch.int32, device=DEV)
            nat.conv1_fwd(x, w, b, pool, am, B, gate_flags=[flags[0:], flags[1:], flags[2:]], want_ptr=want)
        else:
            nat.conv1_fwd(x, w, b, pool, am, B)
        torch.cuda.synchronize()
        outs.append((pool, am))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def _conv2_inputs(B, seed, tf32=False):
    torch.manual_seed(seed)
    q = tf32_trunc if tf32 else (lambda t: bf(t))
    inp = q(torch.rand(B, 14, 14, 32, device=DEV))
    w = q(torch.randn(5, 5, 32, 64, device=DEV) * 0.1)
    b = torch.randn(64, device=DEV) * 0.1
    return inp, w, b


CONV2_ARMS = [("bf16", True), ("bf16", False), ("tf32", False)]  # (precision, CUDA-core validation kernel)


@pytest.mark.parametrize("arm,ref_kernel", CONV2_ARMS)
@pytest.mark.parametrize("B", [128, 6, 75])
def test_conv2_fwd(nat, B, arm, ref_kernel):
    tf32 = arm == "tf32"
    inp, w, b = _conv2_inputs(B, 6, tf32)
    pool = torch.zeros(B, 7, 7, 64, device=DEV, dtype=torch.float32 if tf32 else torch.bfloat16)
    am = torch.zeros(B, 7, 7, 64, device=DEV, dtype=torch.uint8)
    nat.conv2_fwd(inp, w.reshape(800, 64), b, pool, am, B, ref=ref_kernel, tf32=tf32)
    torch.cuda.synchronize()
    pre = torch.relu(oracle.conv2d_same_nhwc(inp.double(), w.double()).float() + b)
    ref = oracle.maxpool2x2_nhwc(pre)
    assert rel_err(pool, ref) < (2e-5 if tf32 else 1e-2)
    # argmax points at a maximal element of its window
    win = pre.reshape(B, 7, 2, 7, 2, 64).permute(0, 1, 3, 5, 2, 4).reshape(B, 7, 7, 64, 4)
    picked = torch.gather(win, 4, am.long().unsqueeze(-1)).squeeze(-1)
    assert (picked - ref).abs().max().item() <= (1e-4 if tf32 else 2e-2) * ref.abs().max().item()


@pytest.mark.parametrize("arm,ref_kernel", CONV2_ARMS)
@pytest.mark.parametrize("B", [128, 6, 75])
def test_conv2_dgrad(nat, B, arm, ref_kernel):
    tf32 = arm == "tf32"
    inp, w, b = _conv2_inputs(B, 7, tf32)
    dy = torch.randn(B, 14, 14, 64, device=DEV)
    dy = tf32_trunc(dy) if tf32 else bf(dy)
    xin = inp.double().requires_grad_(True)
    (oracle.conv2d_same_nhwc(xin, w.double()) * dy.double()).sum().backward()
    nplanes = 2 if tf32 else 1
    dx = torch.full((nplanes, B, 14, 14, 32), 7.0, device=DEV)
    nat.conv2_dgrad(dy, w.reshape(800, 64), dx, B, ref=ref_kernel, tf32=tf32)
    torch.cuda.synchronize()
    assert rel_err(dx.sum(0), xin.grad.float()) < (2e-5 if tf32 else 2e-3)


@pytest.mark.parametrize("arm,ref_kernel", CONV2_ARMS)
@pytest.mark.parametrize("B", [128, 6, 75])
def test_conv2_wgrad(nat, B, arm, ref_kernel):
    tf32 = arm == "tf32"
    inp, w, b = _conv2_inputs(B, 8, tf32)
    dy = torch.randn(B, 14, 14, 64, device=DEV)
    dy = tf32_trunc(dy) if tf32 else bf(dy)
    wl = w.double().requires_grad_(True)
    (oracle.conv2d_same_nhwc(inp.double(), wl) * dy.double()).sum().backward()
    scratch = _scratch(nat, "conv2_wgrad", B)
    res = []
    for _ in range(2):  # written (not accumulated) in a fixed order: two launches on one scratch agree bit for bit
        dw = torch.full((800, 64), 7.0, device=DEV)
        nat.conv2_wgrad(inp, dy, dw, scratch, B, ref=ref_kernel, tf32=tf32)
        torch.cuda.synchronize()
        res.append(dw)
    assert rel_err(res[0], wl.grad.float().reshape(800, 64)) < (2e-5 if tf32 else 2e-3)
    if not ref_kernel:
        assert torch.equal(res[0], res[1])


@pytest.mark.parametrize("tf32", [False, True])
def test_pool2_bwd(nat, tf32):
    torch.manual_seed(9)
    B = 16
    adt = torch.float32 if tf32 else torch.bfloat16
    pre = torch.randn(B, 14, 14, 64, device=DEV)
    act = torch.relu(pre).requires_grad_(True)
    pooled = oracle.maxpool2x2_nhwc(act)
    dpool = torch.randn(B, 7, 7, 64, device=DEV)
    win = act.detach().reshape(B, 7, 2, 7, 2, 64).permute(0, 1, 3, 5, 2, 4).reshape(B, 7, 7, 64, 4)
    am = win.argmax(dim=4).to(torch.uint8)
    pool_bf = pooled.detach().to(adt)
    dy = torch.empty(B, 14, 14, 64, device=DEV, dtype=adt)
    db = torch.full((64,), 7.0, device=DEV)
    parts = torch.stack([0.5 * dpool, 0.25 * dpool, 0.25 * dpool])
    scratch = _scratch(nat, "pool2_bwd", B)
    nat.pool2_bwd(parts, pool_bf, am, dy, db, scratch, B, nsplit=3, split_stride=dpool.numel(), tf32=tf32)
    torch.cuda.synchronize()
    db1 = db.clone()
    nat.pool2_bwd(parts, pool_bf, am, dy, db, scratch, B, nsplit=3, split_stride=dpool.numel(), tf32=tf32)
    torch.cuda.synchronize()
    assert torch.equal(db, db1)  # written in a fixed order, not accumulated
    g = dpool * (pooled.detach() > 0)
    ref = torch.zeros(B, 7, 7, 64, 4, device=DEV).scatter_(4, am.long().unsqueeze(-1), g.unsqueeze(-1))
    ref = ref.reshape(B, 7, 7, 64, 2, 2).permute(0, 1, 4, 2, 5, 3).reshape(B, 14, 14, 64)
    assert rel_err(dy, ref) < (1e-6 if tf32 else 1e-2)
    assert rel_err(db, g.sum(dim=(0, 1, 2))) < 1e-3


# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tf32", [False, True])
@pytest.mark.parametrize("keep_prob", [1.0, 0.75])
def test_head_cnn_mode_matches_oracle(nat, keep_prob, tf32):
    torch.manual_seed(10)
    B, H = 128, 1024
    adt = torch.float32 if tf32 else torch.bfloat16
    pre = torch.randn(B, H, device=DEV) * 3
    bias = torch.randn(H, device=DEV)
    ow = torch.randn(H, 10, device=DEV) * 0.1
    ob = torch.randn(10, device=DEV)
    y = torch.eye(10, device=DEV)[torch.randint(0, 10, (B,), device=DEV)]
    hid = torch.empty(B, H, device=DEV, dtype=adt)
    dhid = torch.empty(B, H, device=DEV, dtype=adt)
    dl = torch.empty(B, 10, device=DEV)
    lr_ = torch.empty(B, device=DEV)
    hit = torch.empty(B, device=DEV, dtype=torch.uint8)
    mask = torch.ones(B, H, device=DEV, dtype=torch.uint8)
    pre2 = torch.stack([0.25 * pre, 0.75 * pre])  # two split-K partial planes
    nat.head_rows(pre2, None, bias, ow, ob, y, B, H, H, keep_prob, 123, 5, True, hid, dhid, dl, lr_, hit, mask,
                  pre_nsplit=2, pre_split_stride=B * H, tf32=tf32)
    dow = torch.empty(H, 10, device=DEV)
    dhb = torch.empty(H, device=DEV)
    dob = torch.empty(10, device=DEV)
    sc = torch.empty(2, device=DEV)
    nat.head_cols(hid, dhid, H, dl, lr_, hit, B, H, dow, dhb, dob, sc, tf32=tf32)
    torch.cuda.synchronize()
    if keep_prob < 1:
        frac = mask.float().mean().item()
        assert abs(frac - keep_prob) < 0.01
    # oracle with the same mask
    prel = pre.clone().requires_grad_(True)
    owl, obl, bl = ow.clone().requires_grad_(True), ob.clone().requires_grad_(True), bias.clone().requires_grad_(True)
    h = oracle.dropout_with_mask(torch.relu(prel + bl), mask.bool(), keep_prob)
    logits = h @ owl + obl
    loss = oracle.softmax_xent_mean(logits, y)
    loss.backward()
    assert abs(sc[0].item() - loss.item()) < 2e-2 * max(1.0, abs(loss.item()))
    assert abs(sc[1].item() - oracle.accuracy(logits, y).item()) < 2.0 / B
    tol = 1e-4 if tf32 else 2e-2  # the tf32 arm keeps fp32 activations: only summation order differs
    assert rel_err(hid, h) < (1e-6 if tf32 else 1e-2)
    assert rel_err(dhid, prel.grad) < tol
    assert rel_err(dow, owl.grad) < tol
    assert rel_err(dob, obl.grad) < (1e-4 if tf32 else 1e-2)
    assert rel_err(dhb, bl.grad) < tol


def test_head_mlp_mode_and_eval(nat):
    torch.manual_seed(11)
    B, H, ld = 64, 100, 104
    act = torch.zeros(B, ld, device=DEV, dtype=torch.bfloat16)
    act[:, :H] = bf(torch.relu(torch.randn(B, H, device=DEV)))
    ow = torch.randn(H, 10, device=DEV)
    ob = torch.randn(10, device=DEV)
    y = torch.eye(10, device=DEV)[torch.randint(0, 10, (B,), device=DEV)]
    dhid = torch.zeros(B, ld, device=DEV, dtype=torch.bfloat16)
    dl, lr_, hit = torch.empty(B, 10, device=DEV), torch.empty(B, device=DEV), torch.empty(B, device=DEV, dtype=torch.uint8)
    nat.head_rows(None, act, None, ow, ob, y, B, H, ld, 1.0, 0, 0, True, None, dhid, dl, lr_, hit)
    dow, dhb, dob, sc = torch.empty(H, 10, device=DEV), torch.empty(H, device=DEV), torch.empty(10, device=DEV), torch.empty(2, device=DEV)
    nat.head_cols(act, dhid, ld, dl, lr_, hit, B, H, dow, dhb, dob, sc)
    torch.cuda.synchronize()
    a = act[:, :H].float().requires_grad_(True)
    owl = ow.clone().requires_grad_(True)
    logits = a @ owl + ob
    loss = oracle.softmax_xent_mean(logits, y)
    loss.backward()
    assert abs(sc[0].item() - loss.item()) < 1e-3 * max(1.0, loss.item())
    assert rel_err(dhid[:, :H], a.grad * (a.detach() > 0)) < 1e-2
    assert rel_err(dow, owl.grad) < 1e-3
    # eval: scalars only
    sc2 = torch.zeros(2, device=DEV)
    nat.head_rows(None, act, None, ow, ob, y, B, H, ld, 1.0, 0, 0, False, None, None, dl, lr_, hit)
    nat.head_cols(None, None, ld, dl, lr_, hit, B, H, None, None, None, sc2)
    torch.cuda.synchronize()
    assert abs(sc2[0].item() - loss.item()) < 1e-3 * max(1.0, loss.item())
    assert abs(sc2[1].item() - oracle.accuracy(logits, y).item()) < 1e-6


def test_colsum_cast_random_normal(nat):
    x = bf(torch.randn(128, 200, device=DEV))
    out = torch.empty(200, device=DEV)
    nat.colsum_bf16(x, 200, 128, 200, out)
    src = torch.randn(7, 100, device=DEV)
    dst = torch.full((7, 104), 3.0, device=DEV, dtype=torch.bfloat16)
    nat.cast_bf16(src, dst, 7, 100, 100, 104)
    r = torch.empty(3274634, device=DEV)
    nat.random_normal(r, 42, 1.0)
    r2 = torch.empty(1000, device=DEV)
    nat.random_normal(r2, 42, 1.0)
    torch.cuda.synchronize()
    assert rel_err(out, x.float().sum(0)) < 1e-3
    assert torch.equal(dst[:, :100], bf(src)) and torch.all(dst[:, 100:] == 0)
    assert abs(r.mean().item()) < 5e-3 and abs(r.std().item() - 1.0) < 5e-3  # tf.random_normal: N(0,1) (m:186-195)
    assert torch.equal(r[:1000], r2)  # deterministic in the seed
    k = ((r - r.mean()) ** 4).mean().item() / r.var().item() ** 2
    assert abs(k - 3.0) < 0.05


def test_adam_apply_matches_tf_form_and_publishes_shadow(nat):
    torch.manual_seed(12)
    from tensorflow_distributed_b200.ops.native import Piece
    n = 64 * 40 + 64 * 2  # piece A [40,64] with a flat shadow; piece B [1,100] padded to 104; tail pad
    w0 = torch.randn(n, device=DEV)
    w, m, v = w0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    shadow = torch.zeros(40 * 64 + 104, device=DEV, dtype=torch.bfloat16)
    pieces = [Piece(0, 40 * 64, 64, 64, 0), Piece(40 * 64, 100, 100, 104, 40 * 64)]
    slots = [torch.randn(n, device=DEV) for _ in range(3)]
    ptrs = [s.data_ptr() for s in slots]
    ctrl = torch.zeros(8, dtype=torch.int32, device=DEV)  # [0]=t [1]=scratch [2]=version [3]=global_step
    ref = {"w": w0.clone().cpu()}
    opt = oracle.TFAdam(0.01)
    for step in range(3):
        nat.adam_apply(w, m, v, shadow, ptrs, 3, 1.0 / 3, n, pieces, 0.01, ctrl[0:], None, ctrl[2:], ctrl[3:])
        opt.apply(ref, {"w": (sum(s for s in slots) / 3).cpu()})
    torch.cuda.synchronize()
    assert torch.allclose(w.cpu(), ref["w"], atol=1e-5, rtol=1e-5)
    assert ctrl[0].item() == 3 and ctrl[2].item() == 3 and ctrl[3].item() == 3
    assert torch.equal(shadow[:40 * 64], bf(w[:40 * 64]))
    assert torch.equal(shadow[40 * 64:40 * 64 + 100], bf(w[40 * 64:40 * 64 + 100]))


def test_push_pull_single_gpu(nat):
    src = [torch.randn(1000, device=DEV), torch.randn(4096 + 3, device=DEV)]
    dst = [torch.zeros_like(s) for s in src]
    ctrl = torch.zeros(4, dtype=torch.int32, device=DEV)
    nat.push_multi([s.data_ptr() for s in src], [d.data_ptr() for d in dst], [s.numel() for s in src],
                   [ctrl[0:].data_ptr(), ctrl[1:].data_ptr()], 0.5, ctrl[2:])
    back = [torch.zeros_like(s) for s in src]
    nat.pull_multi([d.data_ptr() for d in dst], [b.data_ptr() for b in back], [s.numel() for s in src],
                   [ctrl[0:].data_ptr(), None], 1)
    torch.cuda.synchronize()
    for s, d, b in zip(src, dst, back):
        assert torch.allclose(d, 0.5 * s) and torch.equal(d, b)
    assert ctrl[0].item() == 1 and ctrl[1].item() == 1 and ctrl[2].item() == 0


def test_gemm_segmented_operands_single_gpu(nat):
    """The sharded-operand paths of the fused pull/push GEMM, with the shards as separate local allocations."""
    torch.manual_seed(13)
    B_, Kin, H = 128, 3136, 1024
    bounds = [0, 1152, 2176, 3136]  # multiples of 128
    x = bf(torch.randn(B_, Kin, device=DEV))
    W = bf(torch.randn(Kin, H, device=DEV))
    dy = bf(torch.randn(B_, H, device=DEV))
    pieces = [W[bounds[i]:bounds[i + 1]].clone() for i in range(3)]
    flags = torch.tensor([5, 5, 7], dtype=torch.int32, device=DEV)
    fl = [flags[i:].data_ptr() for i in range(3)]
    # forward: K rows sharded, version-gated
    planes = torch.zeros(4, B_, H, device=DEV)
    nat.gemm_bf16_seg(x, 0, [p.data_ptr() for p in pieces], 1, B_, H, Kin, planes, mode=nat.EPI_STORE_F32, lda=Kin,
                      ldb=H, ld_out=H, seg_mode=nat.SEG_B_K, seg_rows=bounds, wait_flags=fl, wait_value=5, bn=64,
                      splits=4, split_stride=B_ * H)
    # dgrad: N rows sharded
    dplanes = torch.zeros(2, B_, Kin, device=DEV)
    nat.gemm_bf16_seg(dy, 0, [p.data_ptr() for p in pieces], 0, B_, Kin, H, dplanes, mode=nat.EPI_STORE_F32, lda=H,
                      ldb=H, ld_out=Kin, seg_mode=nat.SEG_B_N, seg_rows=bounds, wait_flags=fl, wait_value=5, bn=64,
                      splits=2, split_stride=B_ * Kin)
    # wgrad: output rows sharded (fused push into per-shard slots)
    slots = [torch.zeros(bounds[i + 1] - bounds[i], H, device=DEV) for i in range(3)]
    nat.gemm_bf16_seg(x, 1, [dy.data_ptr(), 0, 0], 1, Kin, H, B_, None, mode=nat.EPI_STORE_F32, lda=Kin, ldb=H,
                      ld_out=H, seg_mode=nat.SEG_OUT_M, seg_rows=bounds, out_ptrs=[t.data_ptr() for t in slots], bn=128)
    torch.cuda.synchronize()
    assert rel_err(torch.cat(slots, 0), x.float().t() @ dy.float()) < 2e-3
    assert rel_err(planes.sum(0), x.float() @ W.float()) < 2e-3
    assert rel_err(dplanes.sum(0), dy.float() @ W.float().t()) < 2e-3


def test_gemm_tf32_segmented_operands_single_gpu(nat):
    """The same three sharded paths on the tf32 arm (fp32 shards: what a tf32 worker reads from the owners' masters)."""
    torch.manual_seed(16)
    B_, Kin, H = 128, 3136, 1024
    bounds = [0, 1152, 2176, 3136]
    x, W, dy = torch.randn(B_, Kin, device=DEV), torch.randn(Kin, H, device=DEV), torch.randn(B_, H, device=DEV)
    pieces = [W[bounds[i]:bounds[i + 1]].clone() for i in range(3)]
    planes = torch.zeros(4, B_, H, device=DEV)
    nat.gemm_bf16_seg(x, 0, [p.data_ptr() for p in pieces], 1, B_, H, Kin, planes, mode=nat.EPI_STORE_F32, lda=Kin,
                      ldb=H, ld_out=H, seg_mode=nat.SEG_B_K, seg_rows=bounds, bn=64, splits=4, split_stride=B_ * H,
                      tf32=True)
    dplanes = torch.zeros(2, B_, Kin, device=DEV)
    nat.gemm_bf16_seg(dy, 0, [p.data_ptr() for p in pieces], 0, B_, Kin, H, dplanes, mode=nat.EPI_STORE_F32, lda=H,
                      ldb=H, ld_out=Kin, seg_mode=nat.SEG_B_N, seg_rows=bounds, bn=64, splits=2, split_stride=B_ * Kin,
                      tf32=True)
    slots = [torch.zeros(bounds[i + 1] - bounds[i], H, device=DEV) for i in range(3)]
    nat.gemm_bf16_seg(x, 1, [dy.data_ptr(), 0, 0], 1, Kin, H, B_, None, mode=nat.EPI_STORE_F32, lda=Kin, ldb=H,
                      ld_out=H, seg_mode=nat.SEG_OUT_M, seg_rows=bounds, out_ptrs=[t.data_ptr() for t in slots], bn=128,
                      tf32=True)
    torch.cuda.synchronize()
    xd, Wd, dyd = tf32_trunc(x).double(), tf32_trunc(W).double(), tf32_trunc(dy).double()
    assert rel_err(torch.cat(slots, 0), (xd.t() @ dyd).float()) < 2e-5
    assert rel_err(planes.sum(0), (xd @ Wd).float()) < 2e-5
    assert rel_err(dplanes.sum(0), (dyd @ Wd.t()).float()) < 2e-5


def test_backup_workers_decide_and_masked_apply(nat):
    """R-of-N on the device: the shard averages exactly the slots stamped for the next update and ignores stale ones."""
    from tensorflow_distributed_b200.ops.native import Piece
    torch.manual_seed(15)
    n = 64 * 8
    w0 = torch.randn(n, device=DEV)
    w, m, v = w0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    slots = [torch.randn(n, device=DEV) for _ in range(3)]
    ptrs = [s.data_ptr() for s in slots]
    ctrl = torch.zeros(64, dtype=torch.int32, device=DEV)  # [0]=t [1]=scratch [2]=version [3]=gstep [7]=shutdown [16..]=stamps [24]=mask
    pieces = [Piece(0, n, 64, 0, 0)]
    ref = {"w": w0.clone().cpu()}
    opt = oracle.TFAdam(0.01)
    # update 1: workers 0 and 2 delivered for update 1, worker 1 has not pushed yet
    ctrl[16:19] = torch.tensor([1, 0, 1], dtype=torch.int32)
    nat.adam_apply_backup(w, m, v, None, ptrs, 3, 2, n, pieces, 0.01, ctrl[0:], ctrl[16:], ctrl[24:], ctrl[7:], ctrl[2:], ctrl[3:])
    torch.cuda.synchronize()
    assert ctrl[24].item() == 0b101 and ctrl[0].item() == 1 and ctrl[2].item() == 1
    opt.apply(ref, {"w": ((slots[0] + slots[2]) / 2).cpu()})
    # update 2: worker 2's stamp is stale (still 1): only 0 and 1 count
    ctrl[16:19] = torch.tensor([2, 2, 1], dtype=torch.int32)
    nat.adam_apply_backup(w, m, v, None, ptrs, 3, 2, n, pieces, 0.01, ctrl[0:], ctrl[16:], ctrl[24:], ctrl[7:], ctrl[2:], ctrl[3:])
    torch.cuda.synchronize()
    assert ctrl[24].item() == 0b011 and ctrl[3].item() == 2
    opt.apply(ref, {"w": ((slots[0] + slots[1]) / 2).cpu()})
    assert torch.allclose(w.cpu(), ref["w"], atol=1e-5, rtol=1e-5)
    # released ps: shutdown set, nobody delivered -> empty mask, nothing applied
    ctrl[7] = 1
    before = w.clone()
    nat.adam_apply_backup(w, m, v, None, ptrs, 3, 2, n, pieces, 0.01, ctrl[0:], ctrl[16:], ctrl[24:], ctrl[7:], ctrl[2:], ctrl[3:])
    torch.cuda.synchronize()
    assert ctrl[24].item() == 0 and ctrl[0].item() == 2 and torch.equal(w, before)


def test_adam_apply_pub_signals_publishes_mirrors_and_releases_local_versions(nat):
    """The strict-sync cluster step on ONE GPU: the shard's Adam kernel sends the colocated worker's arrival signal
    itself, waits for it, updates, stores the small variable's new values into every worker's mirror (fp32 and bf16)
    and releases every worker's local version word; a gate on those words then passes.  All "worker heaps" are local
    buffers here, so the driver's single-GPU test run exercises the protocol kernels of the multi-GPU path."""
    from tensorflow_distributed_b200.ops.native import Piece
    torch.manual_seed(17)
    small, big = 40 * 64, 64 * 128              # piece 0: mirrored small variable [40,64]; piece 1: a "direct" matrix
    n = small + big
    w0 = torch.randn(n, device=DEV)
    w, m, v = w0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    shadow = torch.zeros(n, device=DEV, dtype=torch.bfloat16)
    pieces = [Piece(0, small, 64, 64, 0), Piece(small, big, 128, 128, small)]
    slot = torch.randn(n, device=DEV)
    ctrl = torch.zeros(64, dtype=torch.int32, device=DEV)   # [0]=t [1]=blocks [2]=version [3]=gstep [4]=arrive [8]=step
    nworkers = 3
    heaps = [torch.full((65536,), 7, device=DEV, dtype=torch.uint8) for _ in range(nworkers)]  # fake worker heaps
    F32_OFF, B16_OFF, VER_OFF = 1024, 32768, 0
    pub = nat.make_pub([(F32_OFF, B16_OFF, 64), (-1, -1, 0)], [h.data_ptr() for h in heaps],
                       [h.data_ptr() + VER_OFF for h in heaps], signal_arrive=[ctrl[4:].data_ptr()],
                       local_step=ctrl[8:].data_ptr())
    for h in heaps:
        h[VER_OFF:VER_OFF + 4] = 0
    ref = {"w": w0.clone().cpu()}
    opt = oracle.TFAdam(0.01)
    for step in range(2):
        nat.adam_apply(w, m, v, shadow, [slot.data_ptr()], 1, 1.0, n, pieces, 0.01, ctrl[0:], ctrl[4:], ctrl[2:], ctrl[3:],
                       pub=pub)
        opt.apply(ref, {"w": slot.cpu()})
    # the worker side: a gate on the local version words passes for step 2, conv1's folded gate too
    want = torch.tensor([2], dtype=torch.int32, device=DEV)
    nat.gate_flags([h.data_ptr() + VER_OFF for h in heaps], want)
    torch.cuda.synchronize()
    assert torch.allclose(w.cpu(), ref["w"], atol=1e-5, rtol=1e-5)
    assert ctrl[0].item() == 2 and ctrl[2].item() == 2 and ctrl[3].item() == 2
    assert ctrl[4].item() == 2 and ctrl[8].item() == 2     # two arrival signals sent, worker step counter advanced twice
    for h in heaps:
        assert h[VER_OFF:VER_OFF + 4].view(torch.int32).item() == 2
        assert torch.equal(h[F32_OFF:F32_OFF + 4 * small].view(torch.float32), w[:small])
        assert torch.equal(h[B16_OFF:B16_OFF + 2 * small].view(torch.bfloat16), bf(w[:small]))
        assert torch.all(h[F32_OFF + 4 * small:B16_OFF] == 7) and torch.all(h[B16_OFF + 2 * small:] == 7)  # nothing else
    assert torch.equal(shadow[small:], bf(w[small:]))      # the big matrix keeps its local shadow, is not mirrored


def test_signal_arrive_counts_once_per_shard_and_advances_the_step(nat):
    ctrl = torch.zeros(16, dtype=torch.int32, device=DEV)
    for _ in range(3):
        nat.signal_arrive([ctrl[0:].data_ptr(), ctrl[1:].data_ptr(), ctrl[2:].data_ptr()], ctrl[8:])
    torch.cuda.synchronize()
    assert ctrl[:3].tolist() == [3, 3, 3] and ctrl[8].item() == 3
