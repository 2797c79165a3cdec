"""Workload runtime: builds the network, the data pipeline, the loss and the base optimizer, and runs
forward/backward — the role of the reference's ``DLTrainer`` (dl_trainer.py:123-876) with the same public
methods (``train test update_model update_optimizer data_iter get_data_shape load_model_from_file
save_checkpoint adjust_learning_rate ...``) and the same hyper-parameter policy (:197-228, :541-605).

B200-first differences:
* the whole step (zero_grad -> forward -> loss -> backward -> compress/send -> receive -> fused update)
  can be captured once in a CUDA graph and replayed (:class:`GraphedStep`): ResNet-50 at batch 32 is
  launch-bound on a B200 when driven op by op from Python;
* inputs are prefetched to the device on a copy stream from pinned host memory; the loss is read back
  asynchronously one step late instead of a blocking ``loss.item()`` every iteration (:713);
* NHWC convolutions are on, ``cudnn.benchmark`` is on except for the ImageNet-shape models (settings.py: with the
  autotuner the ranks of a multi-GPU job land on different algorithms and run at the pace of the slowest; the
  reference pins benchmark=False / deterministic=True, :20-21); GK_CHANNELS_LAST=0/1, GK_CUDNN_BENCHMARK=0/1;
* datasets fall back to synthetic stand-ins of the same shape when the files are absent (no network).
"""
from __future__ import annotations

import os
import time
from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.utils.data
import torch.utils.data.distributed

from . import models, settings, utils
from .data import synthetic
from .lars import LARS
from .models import lstm as lstmpy
from .settings import logger

_support_datasets = ['imagenet', 'cifar10', 'an4', 'ptb', 'mnist']
_support_dnns = ['resnet50', 'googlenet', 'inceptionv4', 'inceptionv3', 'vgg16i', 'alexnet',
                 'resnet20', 'resnet56', 'resnet110', 'vgg19', 'vgg16',
                 'lstman4',
                 'lstm',
                 'mnistnet', 'fcn5net', 'lenet', 'lr']

NUM_CPU_THREADS = int(os.environ.get("GK_NUM_CPU_THREADS", "0"))


class _nvtx:
    """NVTX range (``GK_NVTX=1``) around a phase of the step; a no-op otherwise and on CPUs."""

    def __init__(self, name):
        self.on = settings.NVTX and torch.cuda.is_available()
        self.name = name

    def __enter__(self):
        if self.on:
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
        return False

MnistNet = models.MnistNet


def create_net(num_classes, dnn='resnet20', **kwargs):
    """Model factory (dl_trainer.py:84-120)."""
    ext = None
    if dnn in ('resnet20', 'resnet56', 'resnet110'):
        net = getattr(models, dnn)(num_classes=num_classes)
    elif dnn == 'resnet50':
        import torchvision
        net = torchvision.models.resnet50(num_classes=num_classes)   # like the reference (:90)
    elif dnn == 'inceptionv4':
        net = models.inceptionv4(num_classes=num_classes)
    elif dnn == 'inceptionv3':
        import torchvision
        net = torchvision.models.inception_v3(num_classes=num_classes, init_weights=True)
    elif dnn == 'vgg16i':
        import torchvision
        net = torchvision.models.vgg16(num_classes=num_classes)
    elif dnn == 'googlenet':
        net = models.googlenet()
    elif dnn == 'mnistnet':
        net = models.MnistNet()
    elif dnn == 'fcn5net':
        net = models.FCN5Net()
    elif dnn == 'lenet':
        net = models.LeNet()
    elif dnn == 'lr':
        net = models.LinearRegression()
    elif dnn in ('vgg16', 'vgg19'):
        net = models.VGG(dnn.upper())
    elif dnn == 'alexnet':
        import torchvision
        net = torchvision.models.alexnet()
    elif dnn == 'lstman4':
        net, ext = models.LSTMAN4(datapath=kwargs.get('datapath'))
    elif dnn == 'lstm':
        net = lstmpy.lstm(vocab_size=kwargs['vocab_size'], batch_size=kwargs['batch_size'],
                          num_steps=kwargs.get('num_steps', 35))
    else:
        errstr = 'Unsupport neural network %s' % dnn
        logger.error(errstr)
        raise ValueError(errstr)
    if not hasattr(net, 'name'):
        net.name = dnn
    return net, ext


class _Prefetcher:
    """Double-buffered host->device input pipeline: batch i+1 is copied from pinned memory on a copy
    stream while batch i trains.  ``release`` marks a slot reusable once the step that read it is
    enqueued; the copy stream waits on that before overwriting the slot."""

    def __init__(self, device):
        self.device = device
        self.stream = torch.cuda.Stream(device) if device.type == 'cuda' else None
        self.slots = [None, None]
        self.ready = [None, None]
        self.free = [None, None]
        self.i = 0
        self.bytes = 0

    def put(self, tensors):
        """Start the copy of a tuple of (pinned) host tensors; returns the slot index."""
        s = self.i & 1
        self.i += 1
        if self.stream is None:
            self.slots[s] = tuple(tensors)
            return s
        with torch.cuda.stream(self.stream):
            if self.free[s] is not None:
                self.stream.wait_event(self.free[s])
            if self.slots[s] is None or any(a.shape != b.shape for a, b in zip(self.slots[s], tensors)):
                self.slots[s] = tuple(torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in tensors)
            for d, h in zip(self.slots[s], tensors):
                d.copy_(h, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
            self.ready[s] = ev
        self.bytes = sum(t.numel() * t.element_size() for t in tensors)
        return s

    def get(self, s):
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_event(self.ready[s])
        return self.slots[s]

    def release(self, s):
        if self.stream is not None and s is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.free[s] = ev


class DLTrainer:

    def __init__(self, rank, size, master='gpu10', dist=True, ngpus=1, batch_size=32,
                 is_weak_scaling=True, data_dir='./data', dataset='cifar10', dnn='resnet20',
                 lr=0.04, nworkers=1, prefix=None, sparsity=0.95, pretrain=None, num_steps=35, tb_writer=None,
                 amp_handle=None, optimizer_name='sgd', synthetic_samples=None, device=None):
        self.size = size
        self.rank = rank
        self.pretrain = pretrain
        self.dataset = dataset
        self.prefix = prefix
        self.num_steps = num_steps
        self.ngpus = ngpus
        self.writer = tb_writer
        # Mixed precision (dl_trainer.py:39-45, 256-259, 707-710: apex O2 with a static loss scale of 128, switched
        # by settings.USE_FP16 or an ``amp_handle``): forward and backward run under torch.autocast; parameters,
        # gradients, residuals and the sparse exchange stay fp32 in the arenas (apex O2 keeps fp32 master weights as
        # well).  With fp16 the loss is multiplied by the static scale before backward and the optimizer divides
        # the received gradient by it (DistributedOptimizer.loss_scale); bf16 needs no scaling.
        self.amp_handle = amp_handle
        self.amp = bool(settings.USE_FP16 or amp_handle is not None)
        self.amp_dtype = torch.bfloat16 if settings.AMP_DTYPE.lower() in ("bf16", "bfloat16") else torch.float16
        self.loss_scale = settings.AMP_LOSS_SCALE if (self.amp and self.amp_dtype == torch.float16) else 1.0
        self.synthetic_samples = synthetic_samples
        self.batch_size = batch_size * self.ngpus if (self.ngpus > 0 and is_weak_scaling) else batch_size
        self.num_batches_per_epoch = -1
        self.num_classes = {'cifar10': 10, 'mnist': 10, 'imagenet': 1000, 'an4': 29, 'ptb': 10}[dataset]
        self.nworkers = nworkers  # just for easy comparison
        self.data_dir = data_dir
        self.is_cuda = self.ngpus > 0 and torch.cuda.is_available()
        if device is not None:
            self.device = torch.device(device)
        else:
            self.device = torch.device('cuda', torch.cuda.current_device()) if self.is_cuda else torch.device('cpu')
        self.is_cuda = self.device.type == 'cuda'
        if self.is_cuda:
            torch.backends.cudnn.benchmark = (settings.CUDNN_BENCHMARK if settings.CUDNN_BENCHMARK is not None
                                              else dataset != 'imagenet')
            if settings.CUDNN_BENCHMARK_LIMIT is not None:      # how many algorithms the autotuner tries (0 = all)
                torch.backends.cudnn.benchmark_limit = settings.CUDNN_BENCHMARK_LIMIT
        self.trainset = self.testset = self.trainloader = self.testloader = self.train_sampler = None
        self.vocab_size = 10000
        self.decoder = None
        if not isinstance(dnn, str):
            self.net = dnn
            self.dnn = dnn.name
            self.ext = None
        else:
            self.dnn = dnn
            if self.dnn == 'lstm':
                if data_dir is not None:
                    self.data_prepare()
                self.net, self.ext = create_net(self.num_classes, self.dnn, vocab_size=self.vocab_size,
                                                batch_size=self.batch_size, num_steps=self.num_steps)
            elif self.dnn == 'lstman4':
                self.net, self.ext = create_net(self.num_classes, self.dnn, datapath=self.data_dir)
                if data_dir is not None:
                    self.data_prepare()
            else:
                if data_dir is not None:
                    self.data_prepare()
                self.net, self.ext = create_net(self.num_classes, self.dnn)
        self.lr = lr
        self.base_lr = self.lr
        # NHWC on CPUs too: oneDNN's convolutions are faster in it (ResNet-20, batch 32, 4 threads: 20.9 -> 17.5 ms)
        self.channels_last = (settings.CHANNELS_LAST and
                              self.dataset in ('imagenet', 'cifar10') and self.dnn not in ('lstm', 'lstman4'))
        self.net.to(self.device)
        if self.channels_last:
            self.net.to(memory_format=torch.channels_last)
            if settings.FUSED_BN and self.is_cuda:
                from .models.fuse import fuse_bn_act
                self.fused_bn_sites = fuse_bn_act(self.net)
        self.accuracy = 0
        self.loss = 0.0
        self.train_iter = 0
        self.recved_counter = 0
        self.master = master
        self.average_iter = 0
        if self.dataset != 'an4':
            self.criterion = nn.CrossEntropyLoss().to(self.device)
        else:
            # the reference uses warpctc (sum over the batch, then /batch at :691); torch's CTC replaces it
            self.criterion = nn.CTCLoss(blank=0, reduction='sum', zero_infinity=True)
        # hyper-parameters per dataset (dl_trainer.py:197-210)
        weight_decay = 1e-4
        self.m = 0.9
        nesterov = False
        if self.dataset == 'an4':
            self.lstman4_lr_epoch_tag = 0
        elif self.dataset == 'ptb':
            self.m = 0
            weight_decay = 0
        elif self.dataset == 'imagenet':
            self.m = 0.875
            weight_decay = 2 * 3.0517578125e-05
        decay, no_decay = [], []
        for name, param in self.net.named_parameters():
            if not param.requires_grad:
                continue
            if len(param.shape) == 1 or 'bn' in name or 'bias' in name:
                no_decay.append(param)
            else:
                decay.append(param)
        parameters = [{'params': no_decay, 'weight_decay': 0.},
                      {'params': decay, 'weight_decay': weight_decay}]
        if optimizer_name == 'lars':
            # wired only through a commented line in the reference (dl_trainer.py:229)
            self.optimizer = LARS(parameters, lr=self.lr, momentum=self.m, weight_decay=weight_decay,
                                  eeta=0.001, max_epoch=140)
        else:
            self.optimizer = torch.optim.SGD(parameters, lr=self.lr, momentum=self.m, weight_decay=weight_decay,
                                             nesterov=nesterov)
        self.train_epoch = 0
        if self.pretrain is not None and os.path.isfile(self.pretrain):
            self.load_model_from_file(self.pretrain)

        self.sparsities, self.compression_ratios, self.communication_sizes = [], [], []
        self.sparsity = sparsity
        self.avg_loss_per_epoch = 0.0
        self.timer = self.forwardtime = self.backwardtime = self.iotime = 0.0
        self.epochs_info = []
        self.train_acc_top1 = []
        self.num_of_updates_during_comm = 0
        self._prefetch = _Prefetcher(self.device)
        self._next_slot = None
        self._pending = []          # lagged device->host reads: (event, pinned tensor)
        self._graph: Optional[GraphedStep] = None
        import weakref
        from .parallel import comm as _comm
        _ref = weakref.ref(self)
        _comm.on_shutdown(lambda: _ref() is not None and _ref().finish())
        self._hidden = None
        self.d2h_bytes = 0
        logger.info('num_batches_per_epoch: %d' % self.num_batches_per_epoch)

    # ------------------------------------------------------------------ small accessors
    def get_acc(self):
        return self.accuracy

    def get_loss(self):
        return self.loss

    def get_model_state(self):
        return self.net.state_dict()

    def get_data_shape(self):
        return self._input_shape, self._output_shape

    def get_train_epoch(self):
        return self.train_epoch

    def get_train_iter(self):
        return self.train_iter

    def set_train_epoch(self, epoch):
        self.train_epoch = epoch

    def set_train_iter(self, iteration):
        self.train_iter = iteration

    def get_num_of_training_samples(self):
        return len(self.trainset)

    def _autocast(self):
        return torch.autocast(self.device.type, dtype=self.amp_dtype, enabled=self.amp)

    def _backward(self, loss):
        (loss * self.loss_scale if self.loss_scale != 1.0 else loss).backward()

    def update_optimizer(self, optimizer):
        self.optimizer = optimizer
        if self.loss_scale != 1.0:
            if not hasattr(optimizer, 'set_loss_scale'):
                raise ValueError('fp16 loss scaling needs the DistributedOptimizer wrapper (GK_AMP_DTYPE=bf16 does not)')
            optimizer.set_loss_scale(self.loss_scale)
        self._graph = None

    def finish(self):
        self.flush_metrics()
        if self._graph is not None:          # captured collectives must be gone before the process group is
            if self.is_cuda:
                torch.cuda.synchronize(self.device)
            self._graph = None
        if self.writer is not None:
            self.writer.close()

    # ------------------------------------------------------------------ checkpoints
    def load_model_from_file(self, filename):
        """Accepts the reference's ``{'iter','epoch','state'}`` files (dl_trainer.py:285-290) and the
        superset this framework writes (optimizer + residual state)."""
        checkpoint = torch.load(filename, map_location=self.device, weights_only=False)
        self.net.load_state_dict(checkpoint['state'])
        self.train_epoch = checkpoint['epoch']
        self.train_iter = checkpoint['iter']
        self._loaded_checkpoint = checkpoint
        logger.info('Load pretrain model: %s, start from epoch %d and iter: %d', filename, self.train_epoch,
                    self.train_iter)

    def restore_optimizer_state(self):
        """Second half of a resume: call after ``update_optimizer`` with the distributed optimizer."""
        ck = getattr(self, '_loaded_checkpoint', None)
        if not ck:
            return
        if 'optimizer' in ck:
            try:
                self.optimizer.load_state_dict(ck['optimizer'])
                rebind = getattr(self.optimizer, '_init_fused_state_from_loaded', None)
                if rebind:
                    rebind()
            except Exception as e:  # noqa: BLE001
                logger.warning('optimizer state not restored: %s', e)
        if 'gk_extra' in ck and hasattr(self.optimizer, 'load_extra_state_dict'):
            self.optimizer.load_extra_state_dict(ck['gk_extra'])

    def checkpoint_path(self, epoch=None):
        epoch = self.train_epoch if epoch is None else epoch
        if self.prefix:
            relative_path = './weights/%s/%s-n%d-bs%d-lr%.4f' % (self.prefix, self.dnn, self.nworkers,
                                                                  self.batch_size, self.base_lr)
        else:
            relative_path = './weights/%s-n%d-bs%d-lr%.4f' % (self.dnn, self.nworkers, self.batch_size, self.base_lr)
        utils.create_path(relative_path)
        return os.path.join(relative_path, '%s-rank%d-epoch%d.pth' % (self.dnn, self.rank, epoch))

    def save_checkpoint(self, state=None, filename=None):
        if state is None:
            state = {'iter': self.train_iter, 'epoch': self.train_epoch, 'state': self.get_model_state()}
            if hasattr(self.optimizer, 'state_dict'):
                state['optimizer'] = self.optimizer.state_dict()
            if hasattr(self.optimizer, 'extra_state_dict'):
                state['gk_extra'] = self.optimizer.extra_state_dict()
        filename = filename or self.checkpoint_path()
        torch.save(state, filename)
        return filename

    def remove_dict(self, dictionary):
        dictionary.clear()

    # ------------------------------------------------------------------ data
    def _have(self, *names):
        return self.data_dir not in (None, 'synthetic') and all(
            os.path.exists(os.path.join(self.data_dir, n)) for n in names)

    def _loaders(self, trainset, testset, test_workers=0, collate=None):
        self.trainset, self.testset = trainset, testset
        train_sampler, shuffle = None, True
        if self.nworkers > 1:
            train_sampler = torch.utils.data.distributed.DistributedSampler(
                trainset, num_replicas=self.nworkers, rank=self.rank)
            train_sampler.set_epoch(0)
            shuffle = False
        self.train_sampler = train_sampler
        if isinstance(trainset, synthetic.SyntheticImageDataset) and self.is_cuda and settings.SYNTHETIC_POOL:
            self.trainloader = synthetic.PinnedBatchPool(trainset, self.batch_size, self.rank, self.nworkers)
        else:
            self.trainloader = torch.utils.data.DataLoader(
                trainset, batch_size=self.batch_size, shuffle=shuffle, num_workers=NUM_CPU_THREADS,
                pin_memory=self.is_cuda, sampler=train_sampler, drop_last=True, collate_fn=collate)
        self.testloader = torch.utils.data.DataLoader(
            testset, batch_size=self.batch_size, shuffle=False, num_workers=test_workers,
            pin_memory=self.is_cuda, collate_fn=collate)

    def _n(self, split):
        n = synthetic.SIZES[self.dataset][0 if split == 'train' else 1]
        if self.synthetic_samples:
            n = min(n, self.synthetic_samples if split == 'train' else max(self.synthetic_samples // 5, self.batch_size))
        return n

    def imagenet_prepare(self):
        image_size = 224
        self._input_shape = (self.batch_size, 3, image_size, image_size)
        self._output_shape = (self.batch_size, 1000)
        hdf5fn = os.path.join(self.data_dir or '', 'imagenet-shuffled.hdf5')
        if self._have('imagenet-shuffled.hdf5'):
            import torchvision.transforms as T
            from .data.hdf5 import DatasetHDF5
            normalize = T.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
            trainset = DatasetHDF5(hdf5fn, 'train', T.Compose([T.ToPILImage(), T.RandomResizedCrop(image_size),
                                                               T.RandomHorizontalFlip(), T.ToTensor(), normalize]))
            testset = DatasetHDF5(hdf5fn, 'val', T.Compose([T.ToPILImage(), T.Resize(256), T.CenterCrop(224),
                                                            T.ToTensor(), normalize]))
        else:
            trainset = synthetic.SyntheticImageDataset(self._n('train'), (3, image_size, image_size), 1000, seed=1,
                                                       pool=max(256, 4 * self.batch_size))
            testset = synthetic.SyntheticImageDataset(self._n('test'), (3, image_size, image_size), 1000, seed=2,
                                                      pool=max(64, self.batch_size))
        self._loaders(trainset, testset, test_workers=0)

    def cifar10_prepare(self):
        image_size = 32
        self._input_shape = (self.batch_size, 3, image_size, image_size)
        self._output_shape = (self.batch_size, 10)
        if self._have('cifar-10-batches-py'):
            import torchvision
            import torchvision.transforms as T
            normalize = T.Normalize(mean=[0.491, 0.482, 0.447], std=[0.247, 0.243, 0.262])
            trainset = torchvision.datasets.CIFAR10(root=self.data_dir, train=True, download=False,
                                                    transform=T.Compose([T.RandomCrop(image_size, padding=4),
                                                                         T.RandomHorizontalFlip(), T.ToTensor(),
                                                                         normalize]))
            testset = torchvision.datasets.CIFAR10(root=self.data_dir, train=False, download=False,
                                                   transform=T.Compose([T.ToTensor(), normalize]))
        else:
            trainset = synthetic.SyntheticImageDataset(self._n('train'), (3, 32, 32), 10, seed=1, pool=4096)
            testset = synthetic.SyntheticImageDataset(self._n('test'), (3, 32, 32), 10, seed=1, pool=4096)
        self._loaders(trainset, testset)
        self.classes = ('plane', 'car', 'bird', 'cat', 'deer', 'dog', 'frog', 'horse', 'ship', 'truck')

    def mnist_prepare(self):
        image_size = 32 if self.dnn == 'lenet' else 28
        self._input_shape = (self.batch_size, 1, image_size, image_size)
        self._output_shape = (self.batch_size, 10)
        if self._have('MNIST'):
            import torchvision
            import torchvision.transforms as T
            trans = ([T.Resize(32)] if self.dnn == 'lenet' else []) + [T.ToTensor(), T.Normalize((0.1307,), (0.3081,))]
            trainset = torchvision.datasets.MNIST(self.data_dir, train=True, download=False, transform=T.Compose(trans))
            testset = torchvision.datasets.MNIST(self.data_dir, train=False, transform=T.Compose(trans))
        else:
            trainset = synthetic.SyntheticImageDataset(self._n('train'), (1, image_size, image_size), 10, seed=1, pool=4096)
            testset = synthetic.SyntheticImageDataset(self._n('test'), (1, image_size, image_size), 10, seed=1, pool=4096)
        self._loaders(trainset, testset)

    def ptb_prepare(self):
        from . import ptb_reader
        self._input_shape = (self.batch_size, self.num_steps)
        self._output_shape = (self.batch_size, self.num_steps)
        if self._have('ptb.train.txt'):
            train_data, valid_data, test_data, word_to_id, id_2_word = ptb_reader.ptb_raw_data(data_path=self.data_dir)
            self.vocab_size = len(word_to_id)
            train_set = ptb_reader.TrainDataset(train_data, self.batch_size, self.num_steps)
            test_set = ptb_reader.TestDataset(valid_data, self.batch_size, self.num_steps)
        else:
            self.vocab_size = 10000
            n_tr = self._n('train') if not self.synthetic_samples else self.synthetic_samples * self.num_steps
            n_te = max(n_tr // 12, (self.batch_size + 1) * self.num_steps + 1)
            train_set = synthetic.SyntheticTokenDataset(n_tr, self.vocab_size, self.batch_size, self.num_steps, seed=1)
            test_set = synthetic.SyntheticTokenDataset(n_te, self.vocab_size, self.batch_size, self.num_steps, seed=1)
        self._loaders(train_set, test_set)
        self.testloader = torch.utils.data.DataLoader(test_set, batch_size=self.batch_size, shuffle=False,
                                                      num_workers=0, drop_last=True)

    def an4_prepare(self):
        from .audio_data.data_loader import (AudioDataLoader, BucketingSampler, DistributedBucketingSampler,
                                             SpectrogramDataset, SyntheticSpeechDataset)
        from .decoder import GreedyDecoder
        audio_conf = self.ext['audio_conf']
        labels = self.ext['labels']
        if self._have('an4_train_manifest.csv', 'an4_val_manifest.csv'):
            trainset = SpectrogramDataset(audio_conf=audio_conf,
                                          manifest_filepath=os.path.join(self.data_dir, 'an4_train_manifest.csv'),
                                          labels=labels, normalize=True, augment=True)
            testset = SpectrogramDataset(audio_conf=audio_conf,
                                         manifest_filepath=os.path.join(self.data_dir, 'an4_val_manifest.csv'),
                                         labels=labels, normalize=True, augment=False)
        else:
            trainset = SyntheticSpeechDataset(self._n('train'), labels, seed=1)
            testset = SyntheticSpeechDataset(self._n('test'), labels, seed=2)
        self.trainset, self.testset = trainset, testset
        if self.nworkers > 1:
            train_sampler = DistributedBucketingSampler(trainset, batch_size=self.batch_size,
                                                        num_replicas=self.nworkers, rank=self.rank)
        else:
            train_sampler = BucketingSampler(trainset, batch_size=self.batch_size)
        self.train_sampler = train_sampler
        self.trainloader = AudioDataLoader(trainset, num_workers=NUM_CPU_THREADS, batch_sampler=train_sampler)
        self.testloader = AudioDataLoader(testset, batch_size=self.batch_size, num_workers=NUM_CPU_THREADS)
        self.decoder = GreedyDecoder(labels)
        self._input_shape = (self.batch_size, 1, 161, 200)
        self._output_shape = (self.batch_size, 29)

    def data_prepare(self):
        prepare = {'imagenet': self.imagenet_prepare, 'cifar10': self.cifar10_prepare, 'mnist': self.mnist_prepare,
                   'an4': self.an4_prepare, 'ptb': self.ptb_prepare}.get(self.dataset)
        if prepare is None:
            errstr = 'Unsupport dataset: %s' % self.dataset
            logger.error(errstr)
            raise ValueError(errstr)
        prepare()
        self.data_iterator = iter(self.trainloader)
        # one epoch definition for loop and trainer: full batches only
        self.num_batches_per_epoch = max(
            self.get_num_of_training_samples() // (self.batch_size * self.nworkers), 1)

    def update_nworker(self, nworkers, new_rank=-1):
        """Re-shard the training set for a new world size (dl_trainer.py:508-529)."""
        if new_rank >= 0:
            rank = new_rank
        else:
            reduced = self.nworkers - nworkers
            rank = self.rank - reduced if (reduced > 0 and self.rank >= reduced) else self.rank
        self.rank, self.nworkers = rank, nworkers
        if self.dnn != 'lstman4':
            epoch = self.train_epoch
            self._loaders(self.trainset, self.testset)
            if self.train_sampler is not None:
                self.train_sampler.set_epoch(epoch)
            self.data_iterator = iter(self.trainloader)
        self.num_batches_per_epoch = max(self.get_num_of_training_samples() // (self.batch_size * self.nworkers), 1)

    def data_iter(self):
        try:
            d = next(self.data_iterator)
        except StopIteration:
            self.data_iterator = iter(self.trainloader)
            d = next(self.data_iterator)
        if d[0].size()[0] != self.batch_size:
            return self.data_iter()
        return d

    # ------------------------------------------------------------------ learning-rate schedules
    def _set_lr(self, lr, optimizer):
        self.lr = lr
        for param_group in optimizer.param_groups:
            param_group['lr'] = lr
        return lr

    def _adjust_learning_rate_lstman4(self, progress, optimizer):
        if self.lstman4_lr_epoch_tag != progress:
            self.lstman4_lr_epoch_tag = progress
            self._set_lr(self.lr / 1.01, optimizer)
        return self.lr

    def _adjust_learning_rate_lstmptb(self, progress, optimizer):
        first, second, third = 23 + 40, 60, 80     # dl_trainer.py:548-563
        scale = 1.0 if progress < first else 0.1 if progress < second else 0.01 if progress < third else 0.001
        return self._set_lr(self.base_lr * scale, optimizer)

    def _adjust_learning_rate_general(self, progress, optimizer):
        warmup = 5
        if settings.WARMUP and progress < warmup:
            warmup_total_iters = self.num_batches_per_epoch * warmup
            min_lr = self.base_lr / warmup_total_iters
            lr_interval = (self.base_lr - min_lr) / warmup_total_iters
            return self._set_lr(min_lr + lr_interval * self.train_iter, optimizer)
        first, second, third = 81, 122, 155
        if self.dataset == 'imagenet':
            first, second, third = 30, 60, 80
        elif self.dataset == 'ptb':
            first, second, third = 24, 60, 80
        scale = 1.0 if progress < first else 0.1 if progress < second else 0.01 if progress < third else 0.001
        return self._set_lr(self.base_lr * scale, optimizer)

    def adjust_learning_rate(self, progress, optimizer):
        if self.dnn == 'lstman4':
            return self._adjust_learning_rate_lstman4(self.train_iter // self.num_batches_per_epoch, optimizer)
        elif self.dnn == 'lstm':
            return self._adjust_learning_rate_lstmptb(progress, optimizer)
        return self._adjust_learning_rate_general(progress, optimizer)

    def cal_accuracy(self, output, target, topk=(1,)):
        """Precision@k over a batch, as device tensors (no host sync; the reference's version at
        dl_trainer.py:611-623 is followed by ``float()``)."""
        with torch.no_grad():
            maxk = max(topk)
            _, pred = output.topk(maxk, 1, True, True)
            correct = pred.eq(target.view(-1, 1))
            return [correct[:, :k].any(dim=1).float().mean().mul_(100.0) for k in topk]

    # ------------------------------------------------------------------ metric read-back
    def _enqueue_metrics(self, loss_t, acc_t=None, status_t=None):
        """Asynchronous read-back of the step's loss / accuracy and of the exchange's status words
        [pairs selected, sent, seq, err, withheld]; consumed one step later, so a receive time-out raises within
        one step instead of going unnoticed until the end of training (the reference has no such check at all)."""
        if status_t is None:
            st = getattr(self.optimizer, 'status_tensor', None)
            status_t = st() if callable(st) else None
        vals = torch.stack([loss_t.detach().float().reshape(()),
                            (acc_t if acc_t is not None else loss_t.new_zeros(())).detach().float().reshape(())])
        if status_t is not None and status_t.numel():
            vals = torch.cat([vals, status_t.to(vals.device)])
        if self.is_cuda:
            host = torch.empty(vals.numel(), dtype=torch.float32, pin_memory=True)
            host.copy_(vals, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            self._pending.append((ev, host, acc_t is not None))
        else:
            self._pending.append((None, vals.clone(), acc_t is not None))
        self.d2h_bytes = 4 * vals.numel()
        while len(self._pending) > 1:
            self._consume_metric(self._pending.pop(0))

    def _consume_metric(self, item):
        ev, host, has_acc = item
        if ev is not None:
            ev.synchronize()
        if host.numel() >= 7 and float(host[5]) != 0.0:      # [loss, acc, total, sent, seq, err, overflow]
            code = int(host[5])
            why = {1: "a peer's pairs did not arrive within GK_P2P_TIMEOUT_S (dead or stalled rank)",
                   2: "a received pair list was not index-sorted", 3: "a grid barrier of the chain timed out"}
            raise RuntimeError("sparse gradient exchange failed on rank %d at iteration %d: %s (device error word %d)"
                               % (self.rank, self.train_iter, why.get(code, "unknown"), code))
        loss_value = float(host[0])
        self.loss = loss_value
        self.avg_loss_per_epoch += loss_value
        if has_acc:
            self.train_acc_top1.append(float(host[1]))

    def flush_metrics(self):
        while self._pending:
            self._consume_metric(self._pending.pop(0))

    # ------------------------------------------------------------------ epoch bookkeeping
    def _maybe_roll_epoch(self):
        if self.train_iter % self.num_batches_per_epoch == 0 and self.train_iter > 0:
            self.flush_metrics()
            self.train_epoch += 1
            logger.info('train iter: %d, num_batches_per_epoch: %d', self.train_iter, self.num_batches_per_epoch)
            logger.info('Epoch %d, avg train acc: %f, lr: %f, avg loss: %f' % (
                self.train_iter // self.num_batches_per_epoch,
                float(np.mean(self.train_acc_top1)) if self.train_acc_top1 else 0.0, self.lr,
                self.avg_loss_per_epoch / self.num_batches_per_epoch))
            if self.rank == 0 and self.writer is not None:
                self.writer.add_scalar('cross_entropy', self.avg_loss_per_epoch / self.num_batches_per_epoch, self.train_epoch)
                if self.train_acc_top1:
                    self.writer.add_scalar('top-1_acc', float(np.mean(self.train_acc_top1)), self.train_epoch)
            if self.rank == 0 and self.testloader is not None:
                self.test(self.train_epoch)
            self.sparsities, self.compression_ratios, self.communication_sizes = [], [], []
            self.train_acc_top1 = []
            self.epochs_info.append(self.avg_loss_per_epoch / self.num_batches_per_epoch)
            self.avg_loss_per_epoch = 0.0
            if self.rank == 0 and self.train_epoch % 2 == 0:      # dl_trainer.py:649-661
                self.save_checkpoint()
            if self.train_sampler is not None and self.nworkers > 1 and hasattr(self.train_sampler, 'set_epoch'):
                self.train_sampler.set_epoch(self.train_epoch)

    # ------------------------------------------------------------------ input staging
    def _host_batch(self, data):
        inputs, labels = data
        if self.dnn == 'lstm':
            inputs = inputs.transpose(0, 1).contiguous()
            labels = labels.transpose(0, 1).contiguous()
        if self.is_cuda and not inputs.is_pinned():
            inputs, labels = inputs.pin_memory(), labels.pin_memory()
        return inputs, labels

    def _stage(self, data=None):
        """Host batch -> device tensors in the layout the network consumes.  With ``data=None`` the batch
        comes from the training loader and the NEXT batch's host->device copy is started right away on the
        copy stream, so it overlaps this step's compute."""
        if self.dataset == 'an4':
            data = data if data is not None else self.data_iter()
            inputs, labels_cpu, input_percentages, target_sizes = data
            input_sizes = input_percentages.mul(int(inputs.size(3))).int()
            self._cur_slot = None
            return inputs.to(self.device, non_blocking=True), labels_cpu, input_sizes, target_sizes
        if data is not None:
            slot = self._prefetch.put(self._host_batch(data))
        else:
            if self._next_slot is None:
                self._next_slot = self._prefetch.put(self._host_batch(self.data_iter()))
            slot = self._next_slot
            self._next_slot = self._prefetch.put(self._host_batch(self.data_iter())) if self.is_cuda else None
        inputs, labels = self._prefetch.get(slot)
        self._cur_slot = slot
        if (self.channels_last and inputs.dim() == 4
                and not (settings.STEM_TC and getattr(self.net, '_gk_stem_reads_nchw', False))):
            # (a fused torchvision ResNet reads the NCHW batch in its own space-to-depth kernel: no conversion pass)
            inputs = inputs.contiguous(memory_format=torch.channels_last)
        return inputs, labels

    def _release_inputs(self):
        self._prefetch.release(getattr(self, '_cur_slot', None))

    def _forward_loss(self, staged, hidden=None):
        """Forward + loss (under autocast when mixed precision is on); returns (loss, outputs-or-None,
        labels-or-None, hidden)."""
        with self._autocast():
            return self._forward_loss_impl(staged, hidden)

    def _forward_loss_impl(self, staged, hidden=None):
        if self.dnn == 'lstman4':
            inputs, labels_cpu, input_sizes, target_sizes = staged
            out, output_sizes = self.net(inputs, input_sizes)
            out = out.transpose(0, 1).float().log_softmax(2)  # TxNxH
            loss = self.criterion(out, labels_cpu, output_sizes, target_sizes) / inputs.size(0)
            return loss, None, None, hidden
        inputs, labels = staged
        if self.dnn == 'lstm':
            hidden = lstmpy.repackage_hidden(hidden)
            outputs, hidden = self.net(inputs, hidden)
            loss = self.criterion(outputs.view(-1, self.net.vocab_size), labels.reshape(-1))
            return loss, None, None, hidden
        outputs = self.net(inputs)
        if isinstance(outputs, tuple):        # torchvision inception_v3 / googlenet aux heads
            outputs = outputs[0]
        return self.criterion(outputs, labels), outputs, labels, hidden

    # ------------------------------------------------------------------ the training step
    def train(self, num_of_iters=1, data=None, hidden=None):
        """Forward + backward of ``num_of_iters`` mini-batches (dl_trainer.py:625-740).  Does not zero the
        gradients and does not update — the loop calls ``optimizer.zero_grad()`` / ``update_model()``."""
        s = time.time()
        for _ in range(num_of_iters):
            self.adjust_learning_rate(self.train_epoch, self.optimizer)
            self._maybe_roll_epoch()
            ss = time.time()
            staged = self._stage(data)
            data = None
            self.iotime += (time.time() - ss)
            sforward = time.time()
            with _nvtx('gk:forward'):
                loss, outputs, labels, hidden = self._forward_loss(staged, hidden)
            self.forwardtime += (time.time() - sforward)
            sbackward = time.time()
            with _nvtx('gk:backward'):
                self._backward(loss)
            acc1 = self.cal_accuracy(outputs, labels, topk=(1,))[0] if outputs is not None else None
            self._enqueue_metrics(loss, acc1)
            self._release_inputs()
            self.backwardtime += (time.time() - sbackward)
            self.train_iter += 1
        self.num_of_updates_during_comm += 1
        self.timer += time.time() - s
        self._log_progress()
        if self.dnn == 'lstm':
            return num_of_iters, hidden
        return num_of_iters

    def _log_progress(self, display=40):
        if self.train_iter % display == 0:
            logger.warning('[%3d][%5d/%5d][rank:%d] loss: %.3f, average forward (%f) and backward (%f) time: %f, iotime: %f ' %
                           (self.train_epoch, self.train_iter, self.num_batches_per_epoch, self.rank, self.loss,
                            self.forwardtime / display, self.backwardtime / display, self.timer / display,
                            self.iotime / display))
            self.timer = self.iotime = self.forwardtime = self.backwardtime = 0.0

    def train_and_update(self, data=None, hidden=None, clip_norm=None):
        """One complete optimisation step — zero_grad, forward, backward, gradient exchange, (clip,) update.
        On a GPU the whole step is captured in a CUDA graph after a few eager warm-up steps and replayed;
        the reference-style ``zero_grad(); train(1); update_model()`` sequence stays available and gives
        the same numbers.  ``data`` = a host batch (copied to the device here) or None (training loader)."""
        self.adjust_learning_rate(self.train_epoch, self.optimizer)
        self._maybe_roll_epoch()
        s = time.time()
        staged = self._stage(data)
        self.iotime += time.time() - s
        hidden = self._run_step(staged, hidden, clip_norm)
        self._release_inputs()
        self.timer += time.time() - s
        self._log_progress()
        return hidden

    def step_on_device(self, staged, hidden=None, clip_norm=None):
        """``train_and_update`` for a batch that already lives on the device in the network's layout."""
        self.adjust_learning_rate(self.train_epoch, self.optimizer)
        return self._run_step(staged, hidden, clip_norm)

    def launches_per_step(self) -> int:
        """Kernels of this framework's own library per training step (counted while the step was last
        issued op by op, i.e. during warm-up or graph capture)."""
        return getattr(self, '_step_launches', 0)

    def _run_step(self, staged, hidden, clip_norm):
        from .ops.chain import LAUNCHES
        # the recurrent models stay eager: the weights sit in cuDNN's flat layout inside the arena (no re-pack),
        # but cuDNN's RNN with inter-layer dropout still touches the legacy stream, which a capture cannot
        # contain ("operation would make the legacy stream depend on a capturing blocking stream"), and
        # lstman4's batches change shape from step to step
        use_graph = (self.is_cuda and settings.CUDA_GRAPH and self.dnn not in ('lstm', 'lstman4')
                     and getattr(self.optimizer, '_fused_kind', None) is not None
                     and getattr(self.optimizer, 'graph_safe', False))
        n0 = LAUNCHES["n"]
        if use_graph:
            if self._graph is not None and self._graph.version != getattr(self.optimizer, 'graph_version', 0):
                self._graph = None             # the optimizer re-created buffers the captured step points at
            if self._graph is None:
                self._graph = GraphedStep(self, clip_norm)
            hidden = self._graph.run(staged, hidden)
        else:
            self.optimizer.zero_grad()
            with _nvtx('gk:forward'):
                loss, outputs, labels, hidden = self._forward_loss(staged, hidden)
            with _nvtx('gk:backward'):
                self._backward(loss)
            acc1 = self.cal_accuracy(outputs, labels, topk=(1,))[0] if outputs is not None else None
            if clip_norm is not None and hasattr(self.optimizer, 'clip_grad_norm'):
                self.optimizer.synchronize()
                self.optimizer.clip_grad_norm(clip_norm)
            elif clip_norm is not None:
                torch.nn.utils.clip_grad_norm_(self.net.parameters(), clip_norm)
            self.update_model()
            self._enqueue_metrics(loss, acc1)
        if LAUNCHES["n"] > n0:
            self._step_launches = LAUNCHES["n"] - n0
        self.train_iter += 1
        return hidden

    def update_model(self):
        self.optimizer.step()

    def zero_grad(self):
        self.optimizer.zero_grad()

    # ------------------------------------------------------------------ evaluation
    def test(self, epoch):
        """Full pass over the test loader: top-1/top-5, perplexity (PTB) or WER (AN4)
        (dl_trainer.py:742-824).  Emits the log line that logs/plot.py parses."""
        self.net.eval()
        test_loss, top1_acc, top5_acc = 0.0, [], []
        total = total_steps = total_iters = 0
        costs = total_wer = 0.0
        with torch.no_grad():
            for data in self.testloader:
                if self.dataset == 'an4':
                    inputs, labels_cpu, input_percentages, target_sizes = data
                    input_sizes = input_percentages.mul(int(inputs.size(3))).int()
                    inputs = inputs.to(self.device)
                    out, output_sizes = self.net(inputs, input_sizes)
                    decoded_output, _ = self.decoder.decode(out.float(), output_sizes)
                    split_targets, offset = [], 0
                    for sz in target_sizes.tolist():
                        split_targets.append(labels_cpu[offset:offset + sz])
                        offset += sz
                    target_strings = self.decoder.convert_to_strings(split_targets)
                    for x in range(len(target_strings)):
                        transcript, reference = decoded_output[x][0], target_strings[x][0]
                        total_wer += self.decoder.wer(transcript, reference) / float(max(len(reference.split()), 1))
                    total += inputs.size(0)
                    total_iters += 1
                    continue
                inputs, labels = data
                if self.dnn == 'lstm':
                    inputs = inputs.transpose(0, 1).contiguous().to(self.device)
                    labels = labels.transpose(0, 1).contiguous().to(self.device)
                    hidden = self.net.init_hidden()
                    outputs, hidden = self.net(inputs, hidden)
                    loss = self.criterion(outputs.view(-1, self.net.vocab_size), labels.reshape(-1))
                    test_loss += float(loss)
                    costs += float(loss) * self.net.num_steps
                    total_steps += self.net.num_steps
                else:
                    inputs, labels = inputs.to(self.device), labels.to(self.device)
                    if self.channels_last and inputs.dim() == 4:
                        inputs = inputs.contiguous(memory_format=torch.channels_last)
                    outputs = self.net(inputs)
                    if isinstance(outputs, tuple):
                        outputs = outputs[0]
                    loss = self.criterion(outputs, labels)
                    acc1, acc5 = self.cal_accuracy(outputs, labels, topk=(1, min(5, outputs.size(1))))
                    top1_acc.append(float(acc1))
                    top5_acc.append(float(acc5))
                    test_loss += float(loss)
                total += labels.size(0)
                total_iters += 1
        test_loss /= max(total_iters, 1)
        if self.dnn not in ['lstm', 'lstman4']:
            acc, acc5 = float(np.mean(top1_acc)), float(np.mean(top5_acc))
        elif self.dnn == 'lstm':
            acc, acc5 = float(np.exp(costs / max(total_steps, 1))), 0.0
        else:
            acc, acc5 = total_wer / max(len(self.testloader.dataset), 1), 0.0
        logger.info('Epoch %d, lr: %f, val loss: %f, val top-1 acc: %f, top-5 acc: %f' % (epoch, self.lr, test_loss, acc, acc5))
        self.accuracy = acc
        self.net.train()
        return acc


class GraphedStep:
    """One training step captured in a CUDA graph: zero_grad -> forward -> loss -> backward -> sparsify +
    send (kernels store into the peers' slots) -> receive (bounded spin) + scatter -> [clip] -> fused update.
    Host-side work per step: copy the batch into the static input tensors, refresh the device-side
    hyper-parameters when lr changed, ``cudaGraphLaunch``.  Everything data-dependent (thresholds, counts,
    sequence numbers) lives in device memory, so replays need no re-capture."""

    WARMUP = 3

    def __init__(self, trainer: DLTrainer, clip_norm=None):
        self.t = trainer
        self.clip_norm = clip_norm
        self.graph = None
        self.calls = 0
        self.static = None
        self.static_hidden = None
        self.out_hidden = None
        self.metrics = None
        self.folded = False
        self.has_status = False
        self.version = getattr(trainer.optimizer, 'graph_version', 0)

    def _eager(self, staged, hidden):
        t = self.t
        opt = t.optimizer
        opt.zero_grad()
        with _nvtx('gk:forward'):
            loss, outputs, labels, hidden = t._forward_loss(staged, hidden)
        with _nvtx('gk:backward'):
            t._backward(loss)
        acc1 = t.cal_accuracy(outputs, labels, topk=(1,))[0] if outputs is not None else loss.new_zeros(())
        self.folded = self.clip_norm is None and opt.can_fold()
        if self.folded:
            opt.folded_update()            # receive folded into the update: no dense gradient
        else:
            opt.synchronize()
            if self.clip_norm is not None:
                opt.clip_grad_norm(self.clip_norm)
            opt.fused_update()
        return loss, acc1, hidden

    def run(self, staged, hidden=None):
        t = self.t
        opt = t.optimizer
        opt._refresh_hyper()
        if self.calls < self.WARMUP:
            loss, acc1, hidden = self._eager(staged, hidden)
            opt._steps_done += 1
            t._enqueue_metrics(loss, acc1 if t.dnn != 'lstm' else None)
            self.calls += 1
            return hidden
        if self.graph is None:
            self.static = tuple(torch.empty_like(x) for x in staged)
            for d, s_ in zip(self.static, staged):
                d.copy_(s_)
            if hidden is not None:
                self.static_hidden = tuple(h.detach().clone() for h in hidden)
            st = getattr(opt, 'status_tensor', None)
            self.has_status = callable(st) and st() is not None
            self.metrics = torch.zeros(7 if self.has_status else 2, dtype=torch.float32, device=t.device)
            torch.cuda.synchronize(t.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                loss, acc1, h_out = self._eager(self.static, self.static_hidden)
                self.metrics[0].copy_(loss.detach())
                self.metrics[1].copy_(acc1.detach())
                if self.has_status:
                    self.metrics[2:7].copy_(opt.status_tensor())
                if h_out is not None:
                    self.out_hidden = tuple(h.detach() for h in h_out)
            self.graph = g
        else:
            for d, s_ in zip(self.static, staged):
                d.copy_(s_, non_blocking=True)
            if hidden is not None:
                for d, s_ in zip(self.static_hidden, hidden):
                    if d.data_ptr() != s_.data_ptr():
                        d.copy_(s_, non_blocking=True)
        self.graph.replay()
        opt._steps_done += 1
        opt.train_iter += 1
        opt._random_seed += 1
        if self.folded:
            opt._grad_state = "zero"           # the folded kernel leaves the gradient arena all zero
        else:
            opt._grad_state = "sparse" if opt._sparse and opt._active else "synced"
        t._enqueue_metrics(self.metrics[0], self.metrics[1] if t.dnn != 'lstm' else None,
                           status_t=self.metrics[2:7] if self.has_status else self.metrics[2:2])
        self.calls += 1
        if self.out_hidden is not None:
            return tuple(h.clone() for h in self.out_hidden)
        return hidden


def train_with_single(dnn, dataset, data_dir, nworkers, lr, batch_size, nsteps_update, max_epochs, num_steps=1):
    """Single-device entry (dl_trainer.py:879-904)."""
    if torch.cuda.is_available():
        torch.cuda.set_device(0)
    trainer = DLTrainer(0, nworkers, dist=False, batch_size=batch_size, is_weak_scaling=True,
                        ngpus=1 if torch.cuda.is_available() else 0, data_dir=data_dir, dataset=dataset, dnn=dnn,
                        lr=lr, nworkers=nworkers, prefix='singlegpu', num_steps=num_steps)
    iters_per_epoch = trainer.get_num_of_training_samples() // (nworkers * batch_size * nsteps_update)
    times = []
    display = 40 if iters_per_epoch > 40 else max(iters_per_epoch - 1, 1)
    for epoch in range(max_epochs):
        hidden = trainer.net.init_hidden() if dnn == 'lstm' else None
        for i in range(iters_per_epoch):
            s = time.time()
            trainer.optimizer.zero_grad()
            for j in range(nsteps_update):
                if dnn == 'lstm':
                    _, hidden = trainer.train(1, hidden=hidden)
                else:
                    trainer.train(1)
            trainer.update_model()
            times.append(time.time() - s)
            if i % display == 0 and i > 0:
                time_per_iter = np.mean(times)
                logger.info('Time per iteration including communication: %f. Speed: %f images/s', time_per_iter,
                            batch_size * nsteps_update / time_per_iter)
                times = []
    trainer.finish()
    return trainer


def main(argv=None):
    import argparse
    import logging
    parser = argparse.ArgumentParser(description="Single trainer")
    parser.add_argument('--batch-size', type=int, default=32)
    parser.add_argument('--nsteps-update', type=int, default=1)
    parser.add_argument('--dataset', type=str, default='imagenet', choices=_support_datasets, help='Specify the dataset for training')
    parser.add_argument('--dnn', type=str, default='resnet50', choices=_support_dnns, help='Specify the neural network for training')
    parser.add_argument('--data-dir', type=str, default='./data', help='Specify the data root path')
    parser.add_argument('--lr', type=float, default=0.1, help='Default learning rate')
    parser.add_argument('--max-epochs', type=int, default=settings.MAX_EPOCHS, help='Default maximum epochs to train')
    parser.add_argument('--num-steps', type=int, default=35)
    args = parser.parse_args(argv)
    batch_size = args.batch_size * args.nsteps_update
    relative_path = './logs/singlegpu-%s/%s-n%d-bs%d-lr%.4f-ns%d' % (settings.PREFIX, args.dnn, 1, batch_size, args.lr, args.nsteps_update)
    utils.create_path(relative_path)
    hdlr = logging.FileHandler(os.path.join(relative_path, settings.hostname + '.log'))
    hdlr.setFormatter(settings.formatter)
    logger.addHandler(hdlr)
    logger.info('Configurations: %s', args)
    train_with_single(args.dnn, args.dataset, args.data_dir, 1, args.lr, args.batch_size, args.nsteps_update,
                      args.max_epochs, args.num_steps)


if __name__ == '__main__':
    main()
