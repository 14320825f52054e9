"""Parameter shape lists of the benchmark configurations (BASELINE.json configs C1-C5,
SURVEY.md §8(d)).  Synthetic gradients of exactly these shapes drive bench.py and the parity tests;
no model code is involved — the hot path only sees the parameter arrays."""


def mlp():
    """C1: Gluon MLP 784-512-512-10 — 6 arrays, 669,706 parameters."""
    return [(512, 784), (512,), (512, 512), (512,), (10, 512), (10,)]


def resnet50():
    """C2: ResNet-50 v1 — 161 fp32 arrays, 25,557,032 parameters (53 conv, 53 BN gamma + 53 BN beta,
    fc weight + bias), in definition order."""
    s = [(64, 3, 7, 7), (64,), (64,)]
    inplanes = 64
    for planes, blocks in ((64, 3), (128, 4), (256, 6), (512, 3)):
        for b in range(blocks):
            s += [(planes, inplanes, 1, 1), (planes,), (planes,),
                  (planes, planes, 3, 3), (planes,), (planes,),
                  (planes * 4, planes, 1, 1), (planes * 4,), (planes * 4,)]
            if b == 0:
                s += [(planes * 4, inplanes, 1, 1), (planes * 4,), (planes * 4,)]
            inplanes = planes * 4
    s += [(1000, 2048), (1000,)]
    return s


def _bert(hidden, layers, ffn, vocab=30522, max_pos=512):
    s = [(vocab, hidden), (max_pos, hidden), (2, hidden), (hidden,), (hidden,)]
    for _ in range(layers):
        for _ in range(4):                      # Q, K, V, O projections
            s += [(hidden, hidden), (hidden,)]
        s += [(hidden,), (hidden,)]             # attention LayerNorm
        s += [(ffn, hidden), (ffn,), (hidden, ffn), (hidden,)]
        s += [(hidden,), (hidden,)]             # output LayerNorm
    s += [(hidden, hidden), (hidden,)]          # pooler
    return s


def bert_base():
    """C3: BERT-base — 199 arrays, 109,482,240 parameters."""
    return _bert(768, 12, 3072)


def bert_large():
    """C4: BERT-large — 391 arrays, 335,141,888 parameters."""
    return _bert(1024, 24, 4096)


ROW_SPARSE_TABLE = (1048576, 512)   # C5: 1M x 512 fp32 embedding table
ROW_SPARSE_IDS = 65536              # rows per device per step, drawn with duplicates


def numel(shape):
    n = 1
    for d in shape:
        n *= d
    return n


CONFIGS = {"mlp": mlp, "resnet50": resnet50, "bert_base": bert_base, "bert_large": bert_large}
