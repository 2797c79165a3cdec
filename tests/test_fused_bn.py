"""Fused BN(+residual)(+ReLU): csrc/gk_bn.cu against F.batch_norm + add + relu in fp32, and the model
rewiring of models/fuse.py (CPU: the rewired model must compute exactly what the original did)."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from gaussiank_sgd_b200 import settings
from gaussiank_sgd_b200.models.fuse import FusedBNReLU, fuse_bn_act
from gaussiank_sgd_b200.ops.bn import bn_act, coop_errors, eligible


# ------------------------------------------------------------------------------------- CPU
def test_bn_act_cpu_fallback_matches_modules():
    torch.manual_seed(0)
    bn = nn.BatchNorm2d(8)
    x, r = torch.randn(4, 8, 5, 5), torch.randn(4, 8, 5, 5)
    ref_bn = copy.deepcopy(bn)
    want = F.relu(ref_bn(x) + r)
    got = bn_act(bn, x, residual=r, relu=True)
    assert torch.equal(got, want)
    assert torch.equal(bn.running_mean, ref_bn.running_mean)
    assert int(bn.num_batches_tracked) == 1
    assert not eligible(bn, x)


def test_fuse_torchvision_resnet_cpu_equivalence():
    tv = pytest.importorskip("torchvision")
    torch.manual_seed(1)
    for ctor in (tv.models.resnet18, tv.models.resnet50):
        net = ctor(num_classes=10)
        ref = copy.deepcopy(net)
        keys = list(net.state_dict().keys())
        sites = fuse_bn_act(net)
        assert sites == sum(isinstance(m, nn.BatchNorm2d) for m in net.modules())
        assert list(net.state_dict().keys()) == keys
        x = torch.randn(2, 3, 64, 64)
        out, want = net(x), ref(x)
        assert torch.equal(out, want)
        out.sum().backward()
        want.sum().backward()
        for (n, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
            assert torch.equal(p.grad, q.grad), n
        net.eval(), ref.eval()
        assert torch.equal(net(x), ref(x))
        clone = copy.deepcopy(net)               # bound forwards follow the copy
        assert torch.equal(clone(x), ref(x))


def test_fuse_sequential_pattern_cpu():
    torch.manual_seed(2)
    from gaussiank_sgd_b200.models import VGG
    net = VGG('VGG16')
    ref = copy.deepcopy(net)
    keys = list(net.state_dict().keys())
    n_bn = sum(isinstance(m, nn.BatchNorm2d) for m in net.modules())
    assert fuse_bn_act(net) == n_bn > 0
    assert list(net.state_dict().keys()) == keys
    assert sum(isinstance(m, FusedBNReLU) for m in net.modules()) == n_bn
    x = torch.randn(2, 3, 32, 32)
    assert torch.equal(net(x), ref(x))
    net.load_state_dict(ref.state_dict())


# ------------------------------------------------------------------------------------- GPU
def _reference(bn, x, res, relu):
    y = F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, True, bn.momentum, bn.eps)
    if res is not None:
        y = y + res
    return F.relu(y) if relu else y


CASES = [  # N, C, H, W
    (32, 64, 28, 28), (8, 256, 14, 14), (4, 96, 9, 9), (2, 12, 5, 7), (16, 2048, 7, 7), (3, 512, 1, 1), (2, 4, 3, 3)]


@pytest.mark.gpu
@pytest.mark.parametrize("shape", CASES)
@pytest.mark.parametrize("relu,with_res", [(True, False), (True, True), (False, False), (False, True)])
@pytest.mark.parametrize("coop", [True, False])
def test_fused_bn_matches_pytorch(shape, relu, with_res, coop, monkeypatch):
    monkeypatch.setattr(settings, "FUSED_BN_COOP", coop)
    torch.manual_seed(sum(shape))
    dev = torch.device("cuda:0")
    N, C, H, W = shape
    bn = nn.BatchNorm2d(C).to(dev)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
        bn.running_mean.normal_()
        bn.running_var.uniform_(0.5, 2.0)
    ref = copy.deepcopy(bn)
    x = (torch.randn(N, C, H, W, device=dev) * 2 + 0.5).contiguous(memory_format=torch.channels_last)
    res = torch.randn(N, C, H, W, device=dev).contiguous(memory_format=torch.channels_last) if with_res else None
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    r1 = res.clone().requires_grad_(True) if with_res else None
    r2 = res.clone().requires_grad_(True) if with_res else None
    assert eligible(bn, x1, r1)
    for rep in range(2):                       # twice: the scratch must come back zeroed
        y = bn_act(bn, x1, residual=r1, relu=relu)
        want = _reference(ref, x2, r2, relu)
        assert y.is_contiguous(memory_format=torch.channels_last)
        torch.testing.assert_close(y, want, rtol=1e-5, atol=2e-5)
        torch.testing.assert_close(bn.running_mean, ref.running_mean, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-5, atol=1e-6)
        ref.num_batches_tracked += 1           # F.batch_norm does not touch it, the module does
        assert int(bn.num_batches_tracked) == int(ref.num_batches_tracked) == rep + 1
        dy = torch.randn_like(y)
        for t in (x1, x2, r1, r2, bn.weight, bn.bias, ref.weight, ref.bias):
            if t is not None:
                t.grad = None
        y.backward(dy)
        want.backward(dy)
        scale = max(1.0, float(dy.abs().max()))
        # an output within rounding of zero may land on either side of the ReLU in the two implementations
        sure = (want.detach().abs() > 1e-5) if relu else torch.ones_like(want, dtype=torch.bool)
        torch.testing.assert_close(x1.grad * sure, x2.grad * sure, rtol=1e-4, atol=2e-5 * scale)
        M = N * H * W
        torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=1e-4, atol=1e-5 * M ** 0.5 * 4)
        torch.testing.assert_close(bn.bias.grad, ref.bias.grad, rtol=1e-4, atol=1e-5 * M ** 0.5 * 4)
        if with_res:
            assert torch.equal(r1.grad * sure, r2.grad * sure)
    if "_gk_bn_scratch" in bn.__dict__:
        sc = bn.__dict__["_gk_bn_scratch"]
        assert float(sc[0].abs().sum()) == 0.0 and int(sc[1]) == 0
    assert coop_errors(bn) == 0
    if coop and C % 32 == 0:
        assert "_gk_bn_coop" in bn.__dict__       # the single-kernel path really ran


@pytest.mark.gpu
def test_fused_bn_ineligible_inputs_fall_back():
    dev = torch.device("cuda:0")
    bn = nn.BatchNorm2d(6).to(dev)                       # C % 4 != 0
    x = torch.randn(4, 6, 5, 5, device=dev).contiguous(memory_format=torch.channels_last)
    assert not eligible(bn, x)
    torch.testing.assert_close(bn_act(copy.deepcopy(bn), x), F.relu(copy.deepcopy(bn)(x)))
    bn8 = nn.BatchNorm2d(8).to(dev)
    xc = torch.randn(4, 8, 5, 5, device=dev)             # NCHW-contiguous
    assert not eligible(bn8, xc)
    bn8.eval()
    xl = xc.contiguous(memory_format=torch.channels_last)
    assert not eligible(bn8, xl)
    torch.testing.assert_close(bn_act(bn8, xl), F.relu(bn8(xl)))


@pytest.mark.gpu
def test_fused_resnet50_step_matches_unfused():
    tv = pytest.importorskip("torchvision")
    torch.manual_seed(3)
    dev = torch.device("cuda:0")
    net = tv.models.resnet50(num_classes=100).to(dev).to(memory_format=torch.channels_last)
    ref = copy.deepcopy(net)
    assert fuse_bn_act(net) == 53
    x = torch.randn(8, 3, 96, 96, device=dev).contiguous(memory_format=torch.channels_last)
    t = torch.randint(0, 100, (8,), device=dev)
    old = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    ref64 = copy.deepcopy(ref).double()          # what both fp32 runs are approximations of
    try:
        l1 = F.cross_entropy(net(x), t)
        l2 = F.cross_entropy(ref(x), t)
        l3 = F.cross_entropy(ref64(x.double()), t)
        l1.backward()
        l2.backward()
        l3.backward()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    assert abs(l1.item() - l3.item()) < 1e-4 * max(1.0, abs(l3.item()))
    ours = theirs = 0.0
    for (n, p), (_, q), (_, d) in zip(net.named_parameters(), ref.named_parameters(), ref64.named_parameters()):
        den = float(d.grad.norm()) + 1e-12
        ours = max(ours, float((p.grad.double() - d.grad).norm()) / den)
        theirs = max(theirs, float((q.grad.double() - d.grad).norm()) / den)
    # fp32 round-off is amplified by 53 BN layers and ReLU masks; the fused path must not be worse than cuDNN's
    assert ours < max(3.0 * theirs, 1e-3), (ours, theirs)
    assert coop_errors(net) == 0
    for (n, b), (_, c) in zip(net.named_buffers(), ref.named_buffers()):
        torch.testing.assert_close(b.float(), c.float(), rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
def test_fused_bn_switch_off():
    dev = torch.device("cuda:0")
    bn = nn.BatchNorm2d(8).to(dev)
    x = torch.randn(4, 8, 5, 5, device=dev).contiguous(memory_format=torch.channels_last)
    old = settings.FUSED_BN
    settings.FUSED_BN = False
    try:
        bn_act(bn, x)
        assert "_gk_bn_scratch" not in bn.__dict__
    finally:
        settings.FUSED_BN = old


@pytest.mark.gpu
def test_fused_bn_in_cuda_graph():
    """The cooperative launches are capturable; replays keep producing what eager produces."""
    torch.manual_seed(5)
    dev = torch.device("cuda:0")
    bn = nn.BatchNorm2d(64).to(dev)
    ref = copy.deepcopy(bn)
    x = torch.randn(8, 64, 14, 14, device=dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn_like(x).requires_grad_(True)
    dy = torch.randn_like(x)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):                              # warm-up: scratch allocation, attribute set-up
            gx0, gr0, gw0, gb0 = torch.autograd.grad(bn_act(bn, x, residual=r), [x, r, bn.weight, bn.bias], dy)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        y = bn_act(bn, x, residual=r)
        gx, gr, gw, gb = torch.autograd.grad(y, [x, r, bn.weight, bn.bias], dy)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    want = F.relu(F.batch_norm(x, ref.running_mean, ref.running_var, ref.weight, ref.bias, True, 0.1, ref.eps) + r)
    torch.testing.assert_close(y, want, rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(gx, gx0, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(gw, gw0, rtol=1e-6, atol=1e-5)
    assert int(bn.num_batches_tracked) == 6
    assert coop_errors(bn) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(4, 64, 32, 32), (2, 64, 17, 23), (3, 32, 8, 6), (32, 64, 112, 112)])
def test_bn_relu_maxpool_matches_pytorch(shape):
    from gaussiank_sgd_b200.ops.bn import bn_relu_maxpool
    torch.manual_seed(7)
    dev = torch.device("cuda:0")
    N, C, H, W = shape
    bn = nn.BatchNorm2d(C).to(dev)
    with torch.no_grad():
        bn.weight.uniform_(-1.0, 1.5)          # negative scales too: the maximum must be taken after BN
        bn.bias.uniform_(-0.5, 0.5)
    ref = copy.deepcopy(bn)
    x = (torch.randn(N, C, H, W, device=dev) * 2 + 0.3).contiguous(memory_format=torch.channels_last)
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    for rep in range(2):
        y = bn_relu_maxpool(bn, x1)
        want = F.max_pool2d(F.relu(ref(x2)), 3, 2, 1)
        assert y.shape == want.shape and y.is_contiguous(memory_format=torch.channels_last)
        torch.testing.assert_close(y, want, rtol=1e-5, atol=2e-5)
        torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-5, atol=1e-6)
        assert int(bn.num_batches_tracked) == rep + 1
        dy = torch.randn_like(y)
        for t in (x1, x2, bn.weight, bn.bias, ref.weight, ref.bias):
            t.grad = None
        y.backward(dy)
        want.backward(dy)
        # a pooled value within rounding of a tie (or of zero) may pick another window position
        close = (x1.grad - x2.grad).abs() <= 1e-4 + 1e-4 * x2.grad.abs()
        assert float(close.float().mean()) > 0.999
        M = N * H * W
        torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=2e-3, atol=2e-4 * M ** 0.5)
        torch.testing.assert_close(bn.bias.grad, ref.bias.grad, rtol=2e-3, atol=2e-4 * M ** 0.5)
    assert "_gk_bn_coop" in bn.__dict__ and coop_errors(bn) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("with_res,relu", [(True, True), (False, True), (True, False)])
def test_bn_act_twice_adds_the_two_gradients(with_res, relu):
    torch.manual_seed(9)
    dev = torch.device("cuda:0")
    bn = nn.BatchNorm2d(64).to(dev)
    ref = copy.deepcopy(bn)
    x = torch.randn(8, 64, 14, 14, device=dev).contiguous(memory_format=torch.channels_last)
    r = torch.randn_like(x) if with_res else None
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    r1 = r.clone().requires_grad_(True) if with_res else None
    r2 = r.clone().requires_grad_(True) if with_res else None
    ya, yb = bn_act(bn, x1, residual=r1, relu=relu, twice=True)
    assert ya.data_ptr() == yb.data_ptr()
    want = bn_act(ref, x2, residual=r2, relu=relu)
    d1, d2 = torch.randn_like(want), torch.randn_like(want)
    (ya * d1).sum().backward(retain_graph=True)          # only one of the two gradients present
    gx_single = x1.grad.clone()
    x1.grad = None
    bn.weight.grad = bn.bias.grad = None
    if with_res:
        r1.grad = None
    ((ya * d1).sum() + (yb * d2).sum()).backward()
    want.backward(d1 + d2)
    torch.testing.assert_close(x1.grad, x2.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=1e-4, atol=1e-4)
    if with_res:
        torch.testing.assert_close(r1.grad, r2.grad, rtol=1e-5, atol=1e-6)
    assert not torch.equal(gx_single, x1.grad)


def test_stem_space_to_depth_rewrite_is_exact():
    """7x7/2 convolution on 3 channels == 4x4/1 convolution on the space-to-depth image (models/fuse.py)."""
    from gaussiank_sgd_b200.models.fuse import _stem_conv_s2d
    torch.manual_seed(4)
    conv = nn.Conv2d(3, 16, 7, 2, 3, bias=False).double()
    ref = copy.deepcopy(conv)
    x = torch.randn(2, 3, 20, 28, dtype=torch.float64).contiguous(memory_format=torch.channels_last)
    y, want = _stem_conv_s2d(conv, x), ref(x)
    assert y.shape == want.shape
    torch.testing.assert_close(y, want, rtol=1e-12, atol=1e-12)
    d = torch.randn_like(want)
    y.backward(d)
    want.backward(d)
    assert conv.weight.grad.shape == (16, 3, 7, 7)
    torch.testing.assert_close(conv.weight.grad, ref.weight.grad, rtol=1e-11, atol=1e-11)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2, 32, 32), (3, 40, 56), (32, 224, 224), (5, 224, 224), (3, 48, 80), (2, 64, 320)])
def test_stem_tensor_core_kernel_matches_fp32_convolution(shape):
    """csrc/gk_stem.cu (TMA + tcgen05.mma kind::tf32, accumulator in TMEM) against an fp32 PyTorch convolution of the
    original 7x7 / stride 2 form, within TF32 tolerance; the weight gradient against the same reference."""
    from gaussiank_sgd_b200 import settings
    from gaussiank_sgd_b200.models.fuse import _stem_conv
    from gaussiank_sgd_b200.ops import stem
    N, H, W = shape
    dev = torch.device("cuda:0")
    torch.manual_seed(N + H)
    conv = nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    ref = copy.deepcopy(conv).double()
    x = torch.randn(N, 3, H, W, device=dev).contiguous(memory_format=torch.channels_last)
    assert settings.STEM_TC and settings.STEM_S2D
    if N == 5:
        x = x.contiguous()                      # a plain NCHW batch, as the loader delivers it: read directly
        assert not x.is_contiguous(memory_format=torch.channels_last)
    before = dict(stem.calls)
    y = _stem_conv(conv, x)
    assert stem.calls["fwd"] == before["fwd"] + 1, "the tensor-core path was not taken"
    want = ref(x.double())
    assert y.shape == want.shape and y.is_contiguous(memory_format=torch.channels_last)
    # TF32 keeps 10 mantissa bits of each operand: |error| <~ 2^-10 * sum |x w| over the 147 taps
    scale = F.conv2d(x.double().abs(), ref.weight.abs(), None, 2, 3)
    err = (y.detach().double() - want.detach()).abs()
    worst = float((err / scale.detach()).max())
    assert worst < 2.0 ** -9, worst
    assert float(err.max()) > 0 or H < 8
    d = torch.randn_like(y)
    y.backward(d)
    want.backward(d.double())
    from gaussiank_sgd_b200.ops import native
    took = native.cuda_lib().gk_stem_wgrad_supported(H // 2, W // 2)      # else cuDNN's weight gradient
    assert took == (shape != (3, 40, 56) and shape != (2, 64, 320)) and stem.calls["wgrad"] == before["wgrad"] + took
    gerr = (conv.weight.grad.double() - ref.weight.grad).abs().max() / ref.weight.grad.abs().max()
    assert float(gerr) < 2e-3, float(gerr)
    torch.cuda.synchronize()
    assert all(v == 0 for v in stem.error_codes().values())


@pytest.mark.gpu
def test_fused_resnet_eval_and_nchw_batches_use_the_tensor_core_stem():
    """Evaluation mode (running statistics, no epilogue sums) and a plain NCHW batch straight from a loader both go
    through the tcgen05 stem of a rewired torchvision ResNet and agree with the untouched model within TF32."""
    tv = pytest.importorskip("torchvision")
    from gaussiank_sgd_b200.ops import stem
    torch.manual_seed(11)
    dev = torch.device("cuda:0")
    net = tv.models.resnet18(num_classes=10).to(dev).to(memory_format=torch.channels_last)
    ref = copy.deepcopy(net)
    fuse_bn_act(net)
    x = torch.randn(4, 3, 64, 64, device=dev)                      # NCHW, as a DataLoader delivers it
    for mode in ("train", "eval"):
        getattr(net, mode)(), getattr(ref, mode)()
        before = stem.calls["fwd"]
        with torch.no_grad():
            got, want = net(x), ref(x.contiguous(memory_format=torch.channels_last))
        assert stem.calls["fwd"] == before + 1, mode
        torch.testing.assert_close(got, want, rtol=2e-2, atol=2e-2)
    for (n, b), (_, c) in zip(net.named_buffers(), ref.named_buffers()):          # both stems are tf32, rounded differently
        torch.testing.assert_close(b.float(), c.float(), rtol=2e-2, atol=2e-3)
    assert all(v == 0 for v in stem.error_codes().values())


@pytest.mark.gpu
def test_stem_helper_kernels_match_their_torch_restatement():
    """k_stem_s2d (NHWC and NCHW input) and k_stem_pack_w against ops/stem.py's plain-torch versions, up to the
    round-to-nearest tf32 conversion both kernels apply."""
    from gaussiank_sgd_b200.ops import native, stem
    lib, dev = native.cuda_lib(), torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(9)

    def tf32(t):        # round to nearest, ties away from zero in magnitude (cvt.rna): add half an ulp of the kept part, truncate
        b = t.contiguous().view(torch.int32)
        return ((b + 0x1000) & ~0x1FFF).view(torch.float32)

    x = torch.randn(3, 3, 20, 36, device=dev)
    want = tf32(stem.space_to_depth_reference(x))
    for nchw, src in ((1, x), (0, x.contiguous(memory_format=torch.channels_last))):
        X = torch.full((3, 13, 21, 12), float("nan"), device=dev)
        native.check(lib.gk_stem_s2d_layout(src.data_ptr(), X.data_ptr(), 3, 20, 36, nchw, st))
        assert torch.equal(X, want), nchw
    w = torch.randn(64, 3, 7, 7, device=dev)
    wpk = torch.empty(24 * 512, device=dev)
    native.check(lib.gk_stem_pack_w(w.data_ptr(), wpk.data_ptr(), st))
    ref = tf32(stem.pack_weight_reference(w))                      # [24, 2, 64, 4]; the kernel groups the rows 8 x 8
    assert torch.equal(wpk.view(24, 2, 64, 4), ref)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(4, 64, 64), (32, 224, 224), (3, 96, 160)])
def test_stem_epilogue_statistics_feed_the_batch_norm(shape, monkeypatch):
    """The stem convolution's epilogue accumulates the per-channel sums of its output; bn1 + ReLU + max pool built on
    them must equal the path that runs its own statistics pass over the same convolution output."""
    from gaussiank_sgd_b200.models.fuse import _stem_conv
    from gaussiank_sgd_b200.ops.bn import bn_relu_maxpool
    N, H, W = shape
    dev = torch.device("cuda:0")
    torch.manual_seed(H)
    conv = nn.Conv2d(3, 64, 7, 2, 3, bias=False).to(dev)
    bn_a, pool = nn.BatchNorm2d(64).to(dev), nn.MaxPool2d(3, 2, 1)
    with torch.no_grad():
        bn_a.weight.uniform_(0.5, 1.5)
        bn_a.bias.uniform_(-0.5, 0.5)
    bn_b = copy.deepcopy(bn_a)
    x = (torch.randn(N, 3, H, W, device=dev) * 2 + 0.3).contiguous(memory_format=torch.channels_last)
    y, sums = _stem_conv(conv, x, want_stats=True)
    assert sums is not None and sums.shape[1] == 128
    M = y.numel() // 64
    yf = y.detach().permute(0, 2, 3, 1).reshape(M, 64).double()
    torch.testing.assert_close(sums[:, :64].double().sum(0), yf.sum(0), rtol=1e-5, atol=1e-3 * M ** 0.5)
    torch.testing.assert_close(sums[:, 64:].double().sum(0), (yf * yf).sum(0), rtol=1e-5, atol=1e-3)
    out_a = bn_relu_maxpool(bn_a, y, pool, presums=sums)
    out_b = bn_relu_maxpool(bn_b, y.detach().requires_grad_(True), pool)
    torch.testing.assert_close(out_a, out_b, rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(bn_a.running_mean, bn_b.running_mean, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(bn_a.running_var, bn_b.running_var, rtol=1e-5, atol=1e-6)
    assert int(bn_a.num_batches_tracked) == int(bn_b.num_batches_tracked) == 1
    d = torch.randn_like(out_a)
    out_a.backward(d)                                   # through the BN into the convolution's weight gradient
    assert conv.weight.grad is not None and bool(torch.isfinite(conv.weight.grad).all())
    assert coop_errors(bn_a) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(32, 256, 28, 28), (8, 2048, 7, 7), (4, 96, 9, 9), (16, 512, 14, 14)])
@pytest.mark.parametrize("twice", [False, True])
def test_two_bns_into_one_activation(shape, twice):
    from gaussiank_sgd_b200.ops.bn import bn2_act
    torch.manual_seed(sum(shape))
    dev = torch.device("cuda:0")
    N, C, H, W = shape
    bna, bnb = nn.BatchNorm2d(C).to(dev), nn.BatchNorm2d(C, momentum=0.2).to(dev)
    with torch.no_grad():
        for bn in (bna, bnb):
            bn.weight.uniform_(0.5, 1.5)
            bn.bias.uniform_(-0.5, 0.5)
    ra, rb = copy.deepcopy(bna), copy.deepcopy(bnb)
    xa = (torch.randn(N, C, H, W, device=dev) * 2 + 0.5).contiguous(memory_format=torch.channels_last)
    xb = (torch.randn(N, C, H, W, device=dev) - 0.3).contiguous(memory_format=torch.channels_last)
    a1, b1 = xa.clone().requires_grad_(True), xb.clone().requires_grad_(True)
    a2, b2 = xa.clone().requires_grad_(True), xb.clone().requires_grad_(True)
    for rep in range(2):
        out = bn2_act(bna, a1, bnb, b1, twice=twice)
        want = F.relu(ra(a2) + rb(b2))
        d1 = torch.randn_like(want)
        for t in (a1, b1, a2, b2, bna.weight, bna.bias, bnb.weight, bnb.bias, ra.weight, ra.bias, rb.weight, rb.bias):
            t.grad = None
        if twice:
            ya, yb = out
            assert ya.data_ptr() == yb.data_ptr()
            d2 = torch.randn_like(want)
            torch.testing.assert_close(ya, want, rtol=1e-5, atol=3e-5)
            ((ya * d1).sum() + (yb * d2).sum()).backward()
            want.backward(d1 + d2)
        else:
            torch.testing.assert_close(out, want, rtol=1e-5, atol=3e-5)
            out.backward(d1)
            want.backward(d1)
        sure = want.detach().abs() > 1e-5
        torch.testing.assert_close(a1.grad * sure, a2.grad * sure, rtol=1e-4, atol=5e-5)
        torch.testing.assert_close(b1.grad * sure, b2.grad * sure, rtol=1e-4, atol=5e-5)
        M = N * H * W
        for p, r in ((bna.weight, ra.weight), (bna.bias, ra.bias), (bnb.weight, rb.weight), (bnb.bias, rb.bias)):
            torch.testing.assert_close(p.grad, r.grad, rtol=1e-4, atol=1e-5 * M ** 0.5 * 6)
        for bn, r in ((bna, ra), (bnb, rb)):
            torch.testing.assert_close(bn.running_mean, r.running_mean, rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(bn.running_var, r.running_var, rtol=1e-5, atol=1e-6)
            assert int(bn.num_batches_tracked) == rep + 1
    assert coop_errors(bna) == 0
    if C % 32 == 0:
        assert any(k[-1] == "pair" for k in bna.__dict__["_gk_bn_coop"])


def test_fuse_marks_which_blocks_hand_their_output_out_twice():
    tv = pytest.importorskip("torchvision")
    net = tv.models.resnet50(num_classes=10)
    fuse_bn_act(net)
    flags = [[bool(getattr(b, "_gk_twice", False)) for b in stage] for stage in (net.layer1, net.layer2, net.layer3, net.layer4)]
    # every block whose output feeds another block (identity shortcut, or the next stage's two convolutions);
    # only the very last block of the network returns a plain tensor
    assert flags == [[True] * 3, [True] * 4, [True] * 6, [True, True, False]]
    x = torch.randn(2, 3, 64, 64)
    out = net(x)                                   # CPU: plain path, tuples pass through the stages
    assert out.shape == (2, 10)
