"""DeepSpeech-2 style acoustic model for AN4 (the reference's models/lstm_models.py:148-287 as configured by
models/lstman4.py:8-33): two masked 2-D convolutions over (frequency, time), five uni-directional LSTM layers
of 800 units with batch norm in between, a look-ahead convolution (context 20), batch norm + Linear(800, 29).
27 569 568 parameters in 40 tensors."""
import math
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

supported_rnns = {'lstm': nn.LSTM, 'rnn': nn.RNN, 'gru': nn.GRU}
supported_rnns_inv = {v: k for k, v in supported_rnns.items()}


class SequenceWise(nn.Module):
    """Apply a module to (T*N, H) and restore (T, N, H)."""

    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, x):
        t, n = x.size(0), x.size(1)
        return self.module(x.reshape(t * n, -1)).view(t, n, -1)


class MaskConv(nn.Module):
    """Run a conv stack over (N, C, F, T) and zero everything beyond each utterance's length after every
    layer, so padding never leaks into batch-norm statistics (models/lstm_models.py:56-72)."""

    def __init__(self, seq_module):
        super().__init__()
        self.seq_module = seq_module

    def forward(self, x, lengths):
        for module in self.seq_module:
            x = module(x)
            t = torch.arange(x.size(3), device=x.device).view(1, 1, 1, -1)
            x = x.masked_fill(t >= lengths.to(x.device).view(-1, 1, 1, 1), 0)
        return x, lengths


class InferenceBatchSoftmax(nn.Module):
    def forward(self, x):
        return x if self.training else F.softmax(x, dim=-1)


class BatchRNN(nn.Module):
    def __init__(self, input_size, hidden_size, rnn_type=nn.LSTM, bidirectional=False, batch_norm=True):
        super().__init__()
        self.bidirectional = bidirectional
        self.batch_norm = SequenceWise(nn.BatchNorm1d(input_size)) if batch_norm else None
        self.rnn = rnn_type(input_size=input_size, hidden_size=hidden_size, bidirectional=bidirectional, bias=True)

    def flatten_parameters(self):
        self.rnn.flatten_parameters()

    def forward(self, x, output_lengths):
        if self.batch_norm is not None:
            x = self.batch_norm(x)
        packed = nn.utils.rnn.pack_padded_sequence(x, output_lengths.cpu(), enforce_sorted=False)
        packed, _ = self.rnn(packed)
        x, _ = nn.utils.rnn.pad_packed_sequence(packed, total_length=x.size(0))
        if self.bidirectional:
            x = x.view(x.size(0), x.size(1), 2, -1).sum(2)
        return x


class Lookahead(nn.Module):
    """Look-ahead convolution for uni-directional models (Wang et al. 2016): every feature looks
    ``context`` frames into the future (models/lstm_models.py:108-145)."""

    def __init__(self, n_features, context):
        super().__init__()
        assert context > 0
        self.n_features, self.context = n_features, context
        self.weight = nn.Parameter(torch.empty(n_features, context + 1))
        stdv = 1.0 / math.sqrt(self.weight.size(1))
        self.weight.data.uniform_(-stdv, stdv)

    def forward(self, x):                       # (T, N, H)
        x = F.pad(x.permute(1, 2, 0), (0, self.context))              # (N, H, T + context)
        x = F.conv1d(x, self.weight.unsqueeze(1), groups=self.n_features)
        return x.permute(2, 0, 1).contiguous()


class DeepSpeech(nn.Module):
    def __init__(self, rnn_type=nn.LSTM, labels="abc", rnn_hidden_size=768, nb_layers=5, audio_conf=None,
                 bidirectional=True, context=20):
        super().__init__()
        self.name = 'lstman4'
        self._version = '0.0.1'
        self._hidden_size, self._hidden_layers, self._rnn_type = rnn_hidden_size, nb_layers, rnn_type
        self._audio_conf = audio_conf or {}
        self._labels = labels
        self._bidirectional = bidirectional
        sample_rate = self._audio_conf.get("sample_rate", 16000)
        window_size = self._audio_conf.get("window_size", 0.02)
        num_classes = len(self._labels)
        self.conv = MaskConv(nn.Sequential(
            nn.Conv2d(1, 32, kernel_size=(41, 11), stride=(2, 2), padding=(20, 5)), nn.BatchNorm2d(32),
            nn.Hardtanh(0, 20, inplace=True),
            nn.Conv2d(32, 32, kernel_size=(21, 11), stride=(2, 1), padding=(10, 5)), nn.BatchNorm2d(32),
            nn.Hardtanh(0, 20, inplace=True)))
        freq = int(math.floor((sample_rate * window_size) / 2) + 1)           # 161 at 16 kHz / 20 ms
        freq = int(math.floor(freq + 2 * 20 - 41) / 2 + 1)
        freq = int(math.floor(freq + 2 * 10 - 21) / 2 + 1)
        rnn_input_size = freq * 32
        rnns = [('0', BatchRNN(rnn_input_size, rnn_hidden_size, rnn_type, bidirectional, batch_norm=False))]
        for i in range(nb_layers - 1):
            rnns.append((str(i + 1), BatchRNN(rnn_hidden_size, rnn_hidden_size, rnn_type, bidirectional)))
        self.rnns = nn.Sequential(OrderedDict(rnns))
        self.lookahead = None if bidirectional else nn.Sequential(Lookahead(rnn_hidden_size, context),
                                                                  nn.Hardtanh(0, 20, inplace=True))
        self.fc = nn.Sequential(SequenceWise(nn.Sequential(nn.BatchNorm1d(rnn_hidden_size),
                                                           nn.Linear(rnn_hidden_size, num_classes, bias=False))))
        self.inference_softmax = InferenceBatchSoftmax()

    def get_seq_lens(self, input_length):
        """Output length of the conv stack along time for every utterance."""
        seq_len = input_length
        for m in self.conv.seq_module:
            if isinstance(m, nn.Conv2d):
                seq_len = torch.div(seq_len + 2 * m.padding[1] - m.dilation[1] * (m.kernel_size[1] - 1) - 1,
                                    m.stride[1], rounding_mode='floor') + 1
        return seq_len.int()

    def forward(self, x, lengths):
        output_lengths = self.get_seq_lens(lengths.cpu().int())
        x, _ = self.conv(x, output_lengths)
        n, c, f, t = x.size()
        x = x.view(n, c * f, t).permute(2, 0, 1).contiguous()        # (T, N, H)
        for rnn in self.rnns:
            x = rnn(x, output_lengths)
        if self.lookahead is not None:
            x = self.lookahead(x)
        x = self.fc(x).transpose(0, 1)                               # (N, T, classes)
        return self.inference_softmax(x), output_lengths

    # ---- the reference's second, unused package format (models/lstm_models.py:240-287)
    @staticmethod
    def serialize(model, optimizer=None, epoch=None, iteration=None, loss_results=None, cer_results=None,
                  wer_results=None, avg_loss=None, meta=None):
        package = {'version': model._version, 'hidden_size': model._hidden_size,
                   'hidden_layers': model._hidden_layers, 'rnn_type': supported_rnns_inv.get(model._rnn_type, 'lstm'),
                   'audio_conf': model._audio_conf, 'labels': model._labels, 'state_dict': model.state_dict(),
                   'bidirectional': model._bidirectional}
        for k, v in (('optim_dict', optimizer.state_dict() if optimizer is not None else None), ('avg_loss', avg_loss),
                     ('epoch', None if epoch is None else epoch + 1), ('iteration', iteration), ('meta', meta)):
            if v is not None:
                package[k] = v
        if loss_results is not None:
            package.update(loss_results=loss_results, cer_results=cer_results, wer_results=wer_results)
        return package

    @classmethod
    def load_model_package(cls, package):
        model = cls(rnn_hidden_size=package['hidden_size'], nb_layers=package['hidden_layers'],
                    labels=package['labels'], audio_conf=package['audio_conf'],
                    rnn_type=supported_rnns[package['rnn_type']], bidirectional=package.get('bidirectional', True))
        model.load_state_dict(package['state_dict'])
        return model

    @staticmethod
    def get_param_size(model):
        return sum(p.numel() for p in model.parameters())
