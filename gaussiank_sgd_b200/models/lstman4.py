"""AN4 speech model factory (models/lstman4.py:8-33 in the reference): DeepSpeech with 5 uni-directional
LSTM layers of 800 units, 16 kHz audio, 20 ms Hamming windows with 10 ms stride, 29 output symbols."""
import json
import os

from .deepspeech import DeepSpeech, supported_rnns

DEFAULT_LABELS = "_'ABCDEFGHIJKLMNOPQRSTUVWXYZ "   # labels.json of the reference: blank, apostrophe, A-Z, space


def create_net(nb_labers=5, labels=None, rnn_type='lstm', bidirectional=False, datapath=None, hidden_size=800,
               hidden_layers=5, sample_rate=16000, window_size=0.02, window_stride=0.01, window='hamming',
               noise_dir=None, noise_prob=0.4, noise_min=0.0, noise_max=0.5):
    if labels is None:
        path = os.path.join(datapath, 'labels.json') if datapath else None
        if path and os.path.isfile(path):
            with open(path) as label_file:
                labels = str(''.join(json.load(label_file)))
        else:
            labels = DEFAULT_LABELS
    audio_conf = dict(sample_rate=sample_rate, window_size=window_size, window_stride=window_stride, window=window,
                      noise_dir=noise_dir, noise_prob=noise_prob, noise_levels=(noise_min, noise_max))
    net = DeepSpeech(rnn_hidden_size=hidden_size, nb_layers=hidden_layers, labels=labels,
                     rnn_type=supported_rnns[rnn_type], audio_conf=audio_conf, bidirectional=bidirectional)
    return net, {'audio_conf': audio_conf, 'labels': labels}
