"""Synthetic parameter-shape lists of the models BASELINE.json names.

Only shapes: the workloads are random-initialised tensors of these shapes, there is no model code.
  resnet50      161 tensors /    25,557,032 params  (torchvision resnet50())
  bert_large    398 tensors /   336,226,108 params  (BertForPreTraining, hidden 1024, 24 layers, vocab 30522)
  llama3_8b     291 tensors / 8,030,261,248 params  (hidden 4096, inter 14336, 32 layers, 8 kv heads, vocab 128256)
"""
import math


def resnet50():
    shapes = [(64, 3, 7, 7), (64,), (64,)]
    inplanes = 64
    for planes, blocks in ((64, 3), (128, 4), (256, 6), (512, 3)):
        for b in range(blocks):
            shapes += [(planes, inplanes, 1, 1), (planes,), (planes,),
                       (planes, planes, 3, 3), (planes,), (planes,),
                       (planes * 4, planes, 1, 1), (planes * 4,), (planes * 4,)]
            if b == 0:  # downsample conv + bn
                shapes += [(planes * 4, inplanes, 1, 1), (planes * 4,), (planes * 4,)]
            inplanes = planes * 4
    shapes += [(1000, 2048), (1000,)]
    return shapes


def bert_large(hidden=1024, layers=24, inter=4096, vocab=30522, max_pos=512, type_vocab=2):
    s = [(vocab, hidden), (max_pos, hidden), (type_vocab, hidden), (hidden,), (hidden,)]
    for _ in range(layers):
        s += [(hidden, hidden), (hidden,)] * 3          # q, k, v
        s += [(hidden, hidden), (hidden,), (hidden,), (hidden,)]   # attn out + LayerNorm
        s += [(inter, hidden), (inter,), (hidden, inter), (hidden,), (hidden,), (hidden,)]  # ffn + LayerNorm
    s += [(hidden, hidden), (hidden,)]                   # pooler
    s += [(vocab,), (hidden, hidden), (hidden,), (hidden,), (hidden,)]  # MLM head (decoder weight tied)
    s += [(2, hidden), (2,)]                             # NSP head
    return s


def llama3_8b(hidden=4096, inter=14336, layers=32, heads=32, kv_heads=8, vocab=128256):
    kv = hidden // heads * kv_heads
    s = [(vocab, hidden)]
    for _ in range(layers):
        s += [(hidden, hidden), (kv, hidden), (kv, hidden), (hidden, hidden),
              (inter, hidden), (inter, hidden), (hidden, inter), (hidden,), (hidden,)]
    s += [(hidden,), (vocab, hidden)]
    return s


def numel(shapes):
    return sum(math.prod(x) for x in shapes)


MODELS = {"resnet50": resnet50, "bert_large": bert_large, "llama3_8b": llama3_8b}
