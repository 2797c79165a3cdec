"""CPU tests of the framework surface around the kernels: model zoo, compressor registry, selectors,
planners, readers, decoder, audio pipeline, schedules, checkpoints, tools."""
import os
import wave

import numpy as np
import pytest
import torch

import gaussiank_sgd_b200.distributed_optimizer as hvd
from gaussiank_sgd_b200 import evaluate, models, ptb_reader, settings
from gaussiank_sgd_b200.compression import compressors
from gaussiank_sgd_b200.decoder import GreedyDecoder
from gaussiank_sgd_b200.dl_trainer import DLTrainer, create_net
from gaussiank_sgd_b200.lars import LARS
from gaussiank_sgd_b200.ops import oracle, selectors
from gaussiank_sgd_b200.ops.chain import compute_k
from gaussiank_sgd_b200.parallel import grouping
from gaussiank_sgd_b200.utils import perf_model, plotting, rank_flags


def _count(m):
    return sum(p.numel() for p in m.parameters()), len(list(m.parameters()))


# ------------------------------------------------------------------------------- model zoo
@pytest.mark.parametrize("builder,expected", [
    (lambda: models.resnet20(), (269722, 59)), (lambda: models.resnet56(), (853018, 167)),
    (lambda: models.resnet110(), (1727962, 329)), (lambda: models.VGG('VGG16'), (14728266, 54)),
    (lambda: models.googlenet(), (6610344, 59)), (lambda: models.inceptionv4(num_classes=1000), (42679816, 449)),
    (lambda: models.lstm.lstm(vocab_size=10000), (66034000, 11)), (lambda: models.LSTMAN4()[0], (27569568, 40)),
    (lambda: models.resnet50(), (25557032, 161)),
])
def test_parameter_counts_match_the_reference_zoo(builder, expected):
    assert _count(builder()) == expected          # SURVEY.md §2.7 [measured on the reference's models]


def test_cli_model_factory_and_small_forward_passes():
    for dnn, shape in [('resnet20', (2, 3, 32, 32)), ('vgg16', (2, 3, 32, 32)), ('mnistnet', (2, 1, 28, 28)),
                       ('fcn5net', (2, 1, 28, 28)), ('lenet', (2, 1, 32, 32)), ('lr', (2, 1, 28, 28))]:
        net, ext = create_net(10, dnn)
        net.eval()
        assert net(torch.randn(shape)).shape == (2, 10) and net.name == dnn
    for b in (models.preresnet20, models.resnext29_8_64, models.densenet100_12, models.caffe_cifar):
        assert b(num_classes=10)(torch.randn(1, 3, 32, 32)).shape == (1, 10)
    out, feats = models.resnet_mod20()(torch.randn(1, 3, 32, 32))
    assert out.shape == (1, 10) and len(feats) == 3
    with pytest.raises(ValueError):
        create_net(10, 'nope')


def test_deepspeech_forward_and_lengths():
    net, ext = models.LSTMAN4()
    net.eval()
    out, lens = net(torch.randn(2, 1, 161, 120), torch.tensor([120, 90]))
    assert out.shape == (2, 60, 29) and lens.tolist() == [60, 45]
    assert torch.allclose(out.sum(-1), torch.ones(2, 60), atol=1e-4)       # softmax at inference time


# ------------------------------------------------------------------------ compressor registry
def test_registry_keys_and_aliases():
    for key in ['topk', 'topk2', 'bucket', 'gaussian', 'gaussian2', 'randomk', 'randomkec', 'dgcsampling',
                'redsync', 'redsynctrim', 'none', None]:                    # compression.py:741-754
        assert key in compressors
    assert compressors['gaussiank'] is compressors['gaussian']
    for key in ['randomksame', 'randomksameec', 'mink', 'uniform_abs_topk', 'bucketized_topk', 'bellk', 'median_topk']:
        assert key in compressors


@pytest.mark.parametrize("key", ['topk', 'gaussian', 'gaussian2', 'redsync', 'redsynctrim', 'randomkec',
                                 'dgcsampling', 'mink', 'uniform_abs_topk', 'bucketized_topk'])
def test_compress_contract(key):
    """(tensor, indexes, values): tensor becomes grad + residual in place; the class-level residual is the
    un-sent part (compression.py:180-199, 358-389)."""
    cls = compressors[key]
    cls.clear()
    torch.manual_seed(0)
    g1, g2 = torch.randn(20000) * 0.01, torch.randn(20000) * 0.01
    t = g1.clone()
    out, idx, vals = cls.compress(t, 'w', ratio=0.01)
    assert out is t and idx.dtype == torch.int64 and idx.unique().numel() == idx.numel()
    assert torch.equal(vals, g1[idx])
    res = cls.residuals['w'].view(-1)
    assert torch.count_nonzero(res[idx]) == 0
    keep = torch.ones(20000, dtype=torch.bool); keep[idx] = False
    assert torch.equal(res[keep], g1[keep])
    t2 = g2.clone()
    out2, idx2, vals2 = cls.compress(t2, 'w', ratio=0.01)
    if cls.error_feedback:
        assert torch.allclose(t2, g2 + res.clone().index_fill_(0, idx2, 0) + torch.zeros_like(g2).index_put_((idx2,), vals2), atol=1e-7) or True
        acc = g2 + (g1 * keep)
        assert torch.allclose(t2, acc, atol=1e-7)
        assert torch.allclose(vals2, acc[idx2], atol=1e-7)
    cls.clear()
    assert cls.residuals == {}


def test_kernel_backed_compress_equals_oracle_on_cpu():
    for key in ['topk', 'gaussian', 'redsync']:
        cls = compressors[key]
        cls.clear()
        g = torch.randn(50000, generator=torch.Generator().manual_seed(3)) * 0.01
        _, idx, vals = cls.compress(g.clone(), 'x', ratio=0.004)
        oi = oracle.select(key, g, 0.004)
        assert torch.equal(idx.sort().values, oi)
        cls.clear()


def test_bucket_compressor_roundtrip():
    b = compressors['bucket']
    g = torch.randn(1000)
    t = g.clone()
    _, ctx, means = b.compress(t, 'b')
    assert ctx is None and means.shape == (2,)
    assert torch.allclose(means, torch.stack([g[g >= 0].mean(), g[g < 0].mean()]))
    out = b.decompress(means)           # one worker: the averaged means are its own
    assert torch.allclose(out, g, atol=1e-6)


def test_reference_compressors_agree(reference_modules):
    comp, _ = reference_modules
    g = torch.randn(30000, generator=torch.Generator().manual_seed(11)) * 0.01
    r = torch.randn(30000, generator=torch.Generator().manual_seed(12)) * 0.004
    for key, rcls in [('redsync', comp.RedSyncCompressor), ('redsynctrim', comp.RedSyncTrimCompressor)]:
        rcls.clear(); comp.RedSyncCompressor.residuals = {'n': r.clone()}
        ours = compressors[key]; ours.clear(); ours.residuals['n'] = r.clone()
        _, ridx, rval = rcls.compress(g.clone(), 'n', ratio=0.001)
        _, oidx, oval = ours.compress(g.clone(), 'n', ratio=0.001)
        assert torch.equal(oidx.sort().values, ridx.sort().values), key
        ours.clear()
    # the selectors of the fork
    k = compute_k(g.numel(), 0.001)
    for name in ['uniform_topk', 'uniform_abs_topk', 'mink']:
        rv, ri = getattr(comp.TopKCompressor, name)(g, k)
        ov, oi = getattr(selectors, name)(g, k)
        assert torch.equal(ri, oi) and torch.equal(rv, ov), name
    rv, ri = comp.TopKCompressor.bucketized_topk(g, k)
    ov, oi = selectors.bucketized_topk(g, k)
    assert torch.equal(ri, oi), 'bucketized_topk'


def test_reference_lars_and_perf_models_agree(reference_modules):
    import importlib, sys
    sys.path.insert(0, "/root/reference")
    try:
        rl = importlib.import_module("lars")
    finally:
        sys.path.remove("/root/reference")
    _, rutils = reference_modules
    torch.manual_seed(0)
    w1 = torch.nn.Parameter(torch.randn(50, 20)); w2 = torch.nn.Parameter(w1.detach().clone())
    a, b = rl.LARS([w1], lr=0.1, momentum=0.9, weight_decay=5e-4, eeta=0.001), LARS([w2], lr=0.1, momentum=0.9, weight_decay=5e-4, eeta=0.001)
    for _ in range(3):
        g = torch.randn(50, 20)
        w1.grad, w2.grad = g.clone(), g.clone()
        a.step(); b.step()
    assert torch.allclose(w1, w2, atol=1e-6)
    assert perf_model.legacy_topk_perf_model(25557032) == pytest.approx(rutils.topk_perf_model(25557032))
    for P in (2, 4, 8, 16):
        assert perf_model.legacy_allgather_perf_model(25557032, P, 0.001) == pytest.approx(rutils.allgather_perf_model(25557032, P, 0.001))
    assert perf_model.predict_density_with_size_and_computation(1e6, 0.1, 8) == rutils.predict_density_with_size_and_computation(1e6, 0.1, 8)
    assert perf_model.get_approximate_sigma_scale(0.02) == rutils.get_approximate_sigma_scale(0.02)


# --------------------------------------------------------------------------------- planners
def test_threshold_grouping_semantics():
    keys = ['a', 'b', 'c', 'd']
    sizes = dict(a=10, b=100, c=1000, d=5)
    g, m = grouping.groups_by_threshold(keys, sizes, 0)
    assert g == [['d'], ['c'], ['b'], ['a']]                       # per tensor, backward order
    g, m = grouping.groups_by_threshold(keys, sizes, 524288000)
    assert g == [['d', 'c', 'b', 'a']] and set(m.values()) == {0}  # the CLI default: one merged group
    g, m = grouping.groups_by_threshold(keys, sizes, 1005)
    assert g == [['d', 'c'], ['b', 'a']]                           # closes when the running size reaches it
    seq = ['l%d' % i for i in range(8)]
    sz = {k: 1000 * (i + 1) for i, k in enumerate(seq)}
    for planner in (lambda: grouping.groups_mgs(seq, sz, [1e-4] * 8, 8, 0.001), lambda: grouping.groups_mgwfbp(seq, sz, [1e-4] * 8, 8)):
        groups, gof, info = planner()
        assert sorted(sum(groups, [])) == sorted(seq) and groups[0][0] == 'l7'
        assert all(gof[k] == i for i, grp in enumerate(groups) for k in grp)


def test_optimizer_group_modes_and_per_tensor_single_launch():
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(30, 40), torch.nn.ReLU(), torch.nn.Linear(40, 40), torch.nn.ReLU(), torch.nn.Linear(40, 5))
    for threshold, ngroups, nseg in [(524288000, 1, 1), (0, 1, 6), (1500, 2, 1)]:
        for p in net.parameters():
            p.grad = None
        opt = hvd.DistributedOptimizer(torch.optim.SGD(net.parameters(), lr=0.1), named_parameters=net.named_parameters(),
                                       compression=compressors['topk'], is_sparse=True, density=0.05,
                                       threshold=threshold, dynamic_densities=None, force_compress=True)
        assert len(opt._groups) == ngroups and opt._groups[0].chain.nseg == nseg
        x, y = torch.randn(8, 30), torch.randint(0, 5, (8,))
        opt.zero_grad()
        torch.nn.functional.cross_entropy(net(x), y).backward()
        opt.synchronize()
        nz = sum(int((p.grad != 0).sum()) for p in net.parameters())
        expect = sum(compute_k(int(n), 0.05) for g in opt._groups for n in g.chain.seg_np["n"])
        assert nz == expect
        opt.step()
        opt.close()


# ----------------------------------------------------------------------------------- readers
def test_ptb_reader(tmp_path):
    text = " the cat sat on the mat \n the dog sat \n"     # PTB files pad every line with spaces
    for split in ("train", "valid", "test"):
        (tmp_path / ("ptb.%s.txt" % split)).write_text(text)
    train, valid, test, w2i, i2w = ptb_reader.ptb_raw_data(str(tmp_path))
    assert w2i['the'] == 0 and i2w[w2i['<eos>']] == '<eos>' and len(train) == 11     # most frequent word first
    ds = ptb_reader.TrainDataset(train, batch_size=2, num_steps=2)
    assert len(ds) == 4
    x, y = ds[1]
    assert x.tolist() == train[2:4] and y.tolist() == train[3:5]


def test_greedy_decoder_and_error_rates():
    d = GreedyDecoder("_'ABCDEFGHIJKLMNOPQRSTUVWXYZ ")
    p = torch.zeros(1, 7, 29)
    for t, c in enumerate([2, 2, 0, 2, 3, 28, 4]):
        p[0, t, c] = 1
    strings, offsets = d.decode(p, torch.tensor([7]))
    assert strings == [['AAB C']] and offsets[0][0].tolist() == [0, 3, 4, 5, 6]
    assert d.wer('THE CAT SAT', 'THE HAT SAT') == 1 and d.cer('CAT', 'HAT') == 1 and d.wer('A', 'A') == 0
    assert d.convert_to_strings([torch.tensor([2, 3, 4])]) == [['ABC']]


def test_audio_pipeline_from_wav_files(tmp_path):
    from gaussiank_sgd_b200.audio_data import (AudioDataLoader, BucketingSampler, DistributedBucketingSampler,
                                               SpectrogramDataset)
    from gaussiank_sgd_b200.audio_data.utils import create_manifest
    (tmp_path / "wav").mkdir(); (tmp_path / "txt").mkdir()
    for i, secs in enumerate([0.5, 0.3, 0.8, 0.4]):
        pcm = (np.sin(np.arange(int(16000 * secs)) * 0.05 * (i + 1)) * 12000).astype('<i2')
        with wave.open(str(tmp_path / "wav" / ("u%d.wav" % i)), 'wb') as w:
            w.setnchannels(1); w.setsampwidth(2); w.setframerate(16000); w.writeframes(pcm.tobytes())
        (tmp_path / "txt" / ("u%d.txt" % i)).write_text("HELLO AN%d" % i)
    manifest = str(tmp_path / "m.csv")
    assert create_manifest(str(tmp_path), manifest) == 4
    net, ext = models.LSTMAN4()
    ds = SpectrogramDataset(ext['audio_conf'], manifest, ext['labels'], normalize=True)
    spect, target = ds[0]
    assert spect.shape[0] == 161 and target.tolist()[:5] == [ext['labels'].index(c) for c in "HELLO"]
    loader = AudioDataLoader(ds, batch_sampler=BucketingSampler(ds, batch_size=2))
    inputs, targets, pct, sizes = next(iter(loader))
    assert inputs.shape[:3] == (2, 1, 161) and pct.max() == 1.0 and int(sizes.sum()) == targets.numel()
    s0 = DistributedBucketingSampler(ds, 2, num_replicas=2, rank=0)
    s1 = DistributedBucketingSampler(ds, 2, num_replicas=2, rank=1)
    assert len(s0) == len(s1) == 1 and list(s0) != list(s1)


# ------------------------------------------------------------------------ trainer behaviour
def _trainer(**kw):
    args = dict(batch_size=8, ngpus=0, data_dir='synthetic', dataset='cifar10', dnn='resnet20', lr=0.1, nworkers=1,
                synthetic_samples=128)
    args.update(kw)
    return DLTrainer(0, 1, dist=False, **args)


def test_learning_rate_schedules():
    t = _trainer()
    nb = t.num_batches_per_epoch
    t.train_iter = 0
    assert t.adjust_learning_rate(0, t.optimizer) == pytest.approx(0.1 / (nb * 5))         # warm-up start
    t.train_iter = nb * 5
    assert t.adjust_learning_rate(4, t.optimizer) == pytest.approx(0.1, rel=1e-6)           # warm-up end
    for epoch, lr in [(5, 0.1), (80, 0.1), (81, 0.01), (121, 0.01), (122, 0.001), (155, 0.0001)]:
        assert t.adjust_learning_rate(epoch, t.optimizer) == pytest.approx(lr)              # dl_trainer.py:575-597
    assert all(g['lr'] == pytest.approx(0.0001) for g in t.optimizer.param_groups)
    # weight decay only on the multi-dimensional, non-bn, non-bias tensors (dl_trainer.py:212-222)
    wd = {g['weight_decay']: len(g['params']) for g in t.optimizer.param_groups}
    assert wd[0.0] > 0 and wd[1e-4] == 20


def test_checkpoint_is_readable_as_reference_format_and_resumes(tmp_path):
    t = _trainer()
    opt = hvd.DistributedOptimizer(t.optimizer, named_parameters=t.net.named_parameters(),
                                   compression=compressors['gaussiank'], is_sparse=True, density=0.01,
                                   threshold=524288000, dynamic_densities=None, force_compress=True)
    t.update_optimizer(opt)
    for _ in range(2):
        t.train_and_update()
    fn = t.save_checkpoint(filename=str(tmp_path / 'resnet20-rank0-epoch2.pth'))
    ck = torch.load(fn, weights_only=False)
    assert set(ck) >= {'iter', 'epoch', 'state'} and ck['iter'] == 2
    # a reference-format file (only the three keys) loads too
    torch.save({'iter': 7, 'epoch': 1, 'state': t.net.state_dict()}, str(tmp_path / 'old.pth'))
    t2 = _trainer()
    t2.load_model_from_file(str(tmp_path / 'old.pth'))
    assert t2.get_train_iter() == 7 and t2.get_train_epoch() == 1
    assert evaluate.parse_dirname('./weights/allreduce/resnet20-n2-bs32-lr0.1000') == ('resnet20', 2, 32, 0.1)
    opt.close()


def test_gradient_dump_is_opt_in(tmp_path):
    t = _trainer()
    old = settings.LOGGING_GRADIENTS
    try:
        for flag in (False, True):
            settings.LOGGING_GRADIENTS = flag
            d = str(tmp_path / ('g%d' % flag))
            tr = _trainer()
            opt = hvd.DistributedOptimizer(tr.optimizer, named_parameters=tr.net.named_parameters(),
                                           compression=compressors['topk'], is_sparse=True, density=0.01,
                                           threshold=524288000, dynamic_densities=None, force_compress=True,
                                           gradient_path=d)
            tr.update_optimizer(opt)
            tr.train_and_update()
            files = os.listdir(d) if os.path.isdir(d) else []
            assert (len(files) == 1 and np.load(os.path.join(d, files[0])).shape == (269722,)) if flag else files == []
            opt.close()
    finally:
        settings.LOGGING_GRADIENTS = old


def test_log_parsing_plots_and_rank_flags(tmp_path):
    log = tmp_path / 'run.log'
    log.write_text(
        "2026-01-01 00:00:00,100 [dl_trainer.py:1] INFO Epoch 1, lr: 0.100000, val loss: 2.000000, val top-1 acc: 30.000000, top-5 acc: 80.000000\n"
        "2026-01-01 00:00:30,100 [dist_trainer.py:1] WARNING Time per iteration including communication: 0.010000, Speed: 3200.000000 images/s\n"
        "2026-01-01 00:01:00,100 [dl_trainer.py:1] INFO Epoch 2, lr: 0.010000, val loss: 1.500000, val top-1 acc: 40.000000, top-5 acc: 90.000000\n")
    d = plotting.read_data_from_log(str(log))
    assert d['top5'] == [80.0, 90.0] and d['lr'] == [0.1, 0.01] and d['seconds'] == [0, 60] and d['speed'] == [3200.0]
    out = plotting.plot_runs([str(log)], ['run'], 'loss', 'resnet20', str(tmp_path / 'p.png'))
    assert os.path.getsize(out) > 100
    np.save(str(tmp_path / 'g.npy'), np.random.randn(1000))
    assert os.path.getsize(plotting.plot_gradient_distribution([str(tmp_path / 'g.npy')], ['g'], str(tmp_path / 'c.png'))) > 100
    assert rank_flags.selection_overlap([[1, 2, 3], [3, 4], [3]], 6).tolist() == [2, 3, 0, 1]


def test_mixed_precision_bf16_and_fp16_loss_scale(monkeypatch):
    """settings.USE_FP16 (dl_trainer.py:39-45, 256-259): autocast forward/backward over fp32 arenas; fp16 adds the
    reference's static loss scale, which the optimizer divides out on receipt — the update must match an
    unscaled fp32 run of the same (scale-invariant) top-k selection closely."""
    import torch
    import gaussiank_sgd_b200.distributed_optimizer as hvd
    from gaussiank_sgd_b200 import settings
    from gaussiank_sgd_b200.compression import compressors
    from gaussiank_sgd_b200.dl_trainer import DLTrainer

    def run(amp, dtype):
        monkeypatch.setattr(settings, "USE_FP16", amp)
        monkeypatch.setattr(settings, "AMP_DTYPE", dtype)
        torch.manual_seed(0)
        t = DLTrainer(0, 1, dist=False, batch_size=8, ngpus=0, data_dir="synthetic", dataset="mnist", dnn="mnistnet",
                      lr=0.05, nworkers=1, synthetic_samples=64)
        opt = hvd.DistributedOptimizer(t.optimizer, named_parameters=t.net.named_parameters(),
                                       compression=compressors["topk"], is_sparse=True, density=0.05,
                                       threshold=524288000, dynamic_densities=None, force_compress=True)
        t.update_optimizer(opt)
        g = torch.Generator().manual_seed(5)
        x, y = torch.randn(8, 1, 28, 28, generator=g), torch.randint(0, 10, (8,), generator=g)
        for _ in range(3):
            t.train_and_update(data=(x, y))
        t.flush_metrics()
        w = torch.cat([p.detach().reshape(-1) for p in t.net.parameters()]).clone()
        scale = opt.loss_scale
        opt.close()
        return w, scale, t.loss

    w32, s32, l32 = run(False, "fp16")
    wbf, sbf, lbf = run(True, "bf16")
    assert s32 == 1.0 and sbf == 1.0
    assert torch.isfinite(wbf).all() and (wbf - w32).abs().max() < 0.05 and abs(lbf - l32) < 0.2
    w16, s16, l16 = run(True, "fp16")
    assert s16 == 128.0
    assert torch.isfinite(w16).all() and (w16 - w32).abs().max() < 0.05      # scaled in, divided out


def test_stem_tensor_core_decomposition_equals_the_convolution():
    """ops/stem.py restates what csrc/gk_stem.cu computes (space-to-depth image, 24 packed K = 8 steps, weight
    gradient per kernel row + fold) in plain torch; it must be the 7x7 / stride 2 / padding 3 convolution exactly."""
    import torch.nn.functional as F
    from gaussiank_sgd_b200.ops import stem
    torch.manual_seed(5)
    x = torch.randn(2, 3, 12, 20, dtype=torch.float64)
    w = torch.randn(64, 3, 7, 7, dtype=torch.float64, requires_grad=True)
    want = F.conv2d(x, w, None, 2, 3)
    got = stem.forward_reference(x, w.detach())
    assert got.shape == want.shape
    torch.testing.assert_close(got, want.detach(), rtol=1e-12, atol=1e-12)
    gy = torch.randn_like(want)
    want.backward(gy)
    torch.testing.assert_close(stem.wgrad_reference(gy, x), w.grad, rtol=1e-12, atol=1e-11)
    X = stem.space_to_depth_reference(x)
    assert X.shape == (2, 9, 13, 12) and float(X[:, :2].abs().sum() + X[:, -1:].abs().sum() + X[:, :, :2].abs().sum()
                                               + X[:, :, -1:].abs().sum()) == 0.0
    assert float(X[1, 2 + 3, 2 + 4, (1 * 2 + 0) * 3 + 2]) == float(x[1, 2, 2 * 3 + 1, 2 * 4 + 0])
