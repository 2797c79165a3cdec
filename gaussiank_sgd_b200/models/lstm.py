"""PTB language model: Embedding(V,1500) -> 2-layer LSTM(1500) with dropout -> Linear(1500,V)
(models/lstm.py:5-47 in the reference; 66 034 000 parameters for V = 10 000).  The embedding and
soft-max matrices (15 M elements each) are what makes this the selection-throughput stress test."""
import torch
import torch.nn as nn


class lstm(nn.Module):
    def __init__(self, vocab_size, embedding_dim=1500, num_steps=35, batch_size=20, num_layers=2,
                 dp_keep_prob=0.35):
        super().__init__()
        self.name = 'lstm'
        self.embedding_dim, self.num_steps, self.batch_size = embedding_dim, num_steps, batch_size
        self.vocab_size, self.dp_keep_prob, self.num_layers = vocab_size, dp_keep_prob, num_layers
        self.dropout = nn.Dropout(1 - dp_keep_prob)
        self.word_embeddings = nn.Embedding(vocab_size, embedding_dim)
        self.lstm = nn.LSTM(embedding_dim, embedding_dim, num_layers=num_layers, dropout=1 - dp_keep_prob)
        self.sm_fc = nn.Linear(embedding_dim, vocab_size)
        self.init_weights()

    def init_weights(self, init_range: float = 0.1):
        with torch.no_grad():
            self.word_embeddings.weight.uniform_(-init_range, init_range)
            self.sm_fc.weight.uniform_(-init_range, init_range)
            self.sm_fc.bias.zero_()

    def init_hidden(self):
        w = next(self.parameters())
        shape = (self.num_layers, self.batch_size, self.embedding_dim)
        return (w.new_zeros(shape), w.new_zeros(shape))

    def forward(self, inputs, hidden):
        emb = self.dropout(self.word_embeddings(inputs))
        out, hidden = self.lstm(emb, hidden)
        logits = self.sm_fc(self.dropout(out).reshape(-1, self.embedding_dim))
        return logits.view(inputs.size(0), inputs.size(1), self.vocab_size), hidden


def repackage_hidden(h):
    """Detach the carried hidden state from its history (truncated BPTT, models/lstm.py:42-47)."""
    if isinstance(h, torch.Tensor):
        return h.detach()
    return tuple(repackage_hidden(v) for v in h)
