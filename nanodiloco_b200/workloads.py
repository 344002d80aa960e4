"""Parameter-shape tables of the Llama sizes BASELINE.json names (shapes only, no arithmetic).

HF `LlamaForCausalLM(LlamaConfig(hidden_size, intermediate_size, num_hidden_layers, ...))` in
`model.parameters()` order with the default vocabulary (32000) and untied embeddings -- what
`nanodiloco/main.py:97-98` builds from `configs/llama_default.json` (tiny 57 tensors /
9,766,528 parameters; 150M-class 111 / 214,983,680; 1B-class 201 / 1,261,529,088).  Used by bench.py and
tools/ to lay out synthetic flat parameter spaces without instantiating the model.
"""
from __future__ import annotations

from typing import List

CONFIGS = {
    "tiny": dict(hidden=128, inter=512, layers=6),              # configs/llama_default.json
    "llama150m": dict(hidden=1024, inter=2688, layers=12),
    "llama1b": dict(hidden=2048, inter=5632, layers=22),
}


def llama_shapes(hidden: int, inter: int, layers: int, vocab: int = 32000) -> List[tuple]:
    shapes = [(vocab, hidden)]
    for _ in range(layers):
        shapes += [(hidden, hidden)] * 4                                # q, k, v, o
        shapes += [(inter, hidden), (inter, hidden), (hidden, inter)]   # gate, up, down
        shapes += [(hidden,), (hidden,)]                                # input / post-attention norms
    shapes += [(hidden,), (vocab, hidden)]                              # final norm, lm_head
    return shapes


def workload_shapes(name: str) -> List[tuple]:
    return llama_shapes(**CONFIGS[name])
