"""BERT-base encoder (12 layers, hidden 768, 12 heads, FFN 3072) -- pipeline workload of BASELINE.json config 4.

NOT IN REFERENCE (the reference has no attention model).  On CUDA every op of a layer is one of this library's kernels:
projection / FFN matmuls and the six matrix products of attention are tcgen05 GEMMs (``FusedLinear``, ``OF.attention``),
LayerNorm (+ residual), GELU and the row softmax are the kernels of csrc/kernels/elementwise.cu; CPU tensors (and
``use_reference_ops``) run the plain PyTorch restatement.  ``as_sequential`` exposes embeddings, the 12 layers and the head
as a flat list for the layer-split partitioner; activations between stages are a single [B, S, H] tensor.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import functional as OF
from .layers import FusedLinear


class BertEmbeddings(nn.Module):
    def __init__(self, vocab, hidden, max_pos, type_vocab=2, dropout=0.0):
        super().__init__()
        self.word_embeddings = nn.Embedding(vocab, hidden)
        self.position_embeddings = nn.Embedding(max_pos, hidden)
        self.token_type_embeddings = nn.Embedding(type_vocab, hidden)
        self.LayerNorm = nn.LayerNorm(hidden, eps=1e-12)
        self.dropout = nn.Dropout(dropout)

    def forward(self, input_ids):
        S = input_ids.shape[1]
        pos = torch.arange(S, device=input_ids.device)
        x = self.word_embeddings(input_ids) + self.position_embeddings(pos)[None] + self.token_type_embeddings.weight[0]
        x = self.dropout(OF.layer_norm(x, self.LayerNorm))
        return x.to(torch.bfloat16) if OF._use_kernels(x) else x


class BertLayer(nn.Module):
    def __init__(self, hidden=768, heads=12, ffn=3072, dropout=0.0):
        super().__init__()
        self.heads = heads
        self.qkv = FusedLinear(hidden, 3 * hidden)
        self.attn_out = FusedLinear(hidden, hidden)
        self.ln1 = nn.LayerNorm(hidden, eps=1e-12)
        self.ffn_in = FusedLinear(hidden, ffn)
        self.ffn_out = FusedLinear(ffn, hidden)
        self.ln2 = nn.LayerNorm(hidden, eps=1e-12)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x):
        a = OF.attention(self.qkv(x), self.heads)
        x = OF.layer_norm(self.dropout(self.attn_out(a)), self.ln1, residual=x)
        h = self.ffn_out(OF.gelu(self.ffn_in(x)))
        return OF.layer_norm(self.dropout(h), self.ln2, residual=x)


class BertHead(nn.Module):
    """[CLS] pooler + classifier (sequence classification)."""

    def __init__(self, hidden, num_classes):
        super().__init__()
        self.pooler = FusedLinear(hidden, hidden)
        self.classifier = FusedLinear(hidden, num_classes)

    def forward(self, x):
        return self.classifier(torch.tanh(self.pooler(x[:, 0])))


class BertBase(nn.Module):
    def __init__(self, vocab=30522, hidden=768, layers=12, heads=12, ffn=3072, max_pos=512, num_classes=2):
        super().__init__()
        self.embeddings = BertEmbeddings(vocab, hidden, max_pos)
        self.encoder = nn.Sequential(*[BertLayer(hidden, heads, ffn) for _ in range(layers)])
        self.head = BertHead(hidden, num_classes)

    def forward(self, input_ids):
        return self.head(self.encoder(self.embeddings(input_ids)))

    def as_sequential(self) -> nn.Sequential:
        return nn.Sequential(self.embeddings, *list(self.encoder), self.head)
