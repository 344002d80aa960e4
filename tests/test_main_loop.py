"""The training loop that calls the hot path (nanodiloco_b200/main.py, drop-in for the reference's
nanodiloco/main.py:41-130).  CPU: loop bookkeeping with the reference port standing in for the CUDA
wrapper.  GPU: the same loop with nanodiloco_b200.Diloco against the port run on the same GPU."""
import numpy as np
import pytest
import torch

from nanodiloco_b200.main import _parse, train_model
from oracle.torch_reference import ReferenceDilocoModelPort

TINY = {"architectures": ["LlamaForCausalLM"], "hidden_size": 32, "intermediate_size": 64,
        "num_attention_heads": 2, "num_hidden_layers": 2, "rms_norm_eps": 1e-05, "use_cache": False,
        "vocab_size": 128}
KW = dict(batch_size=4, per_device_batch_size=2, seq_length=16, warmup_steps=2, total_steps=8, inner_steps=2,
          max_steps=6, llama_config=TINY, log_every=1)


def test_cli_flags_and_defaults_match_the_reference():
    a = _parse([])
    # main.py:42-56
    assert (a.seed, a.batch_size, a.per_device_batch_size, a.seq_length) == (1337, 256, 8, 1024)
    assert (a.warmup_steps, a.total_steps, a.inner_steps, a.lr, a.outer_lr) == (100, 10_000, 100, 4e-4, 0.7)
    assert a.project == "nano-diloco" and a.llama_config_file is None and a.wandb_config_file is None
    b = _parse(["--inner-steps", "500", "--outer-lr", "0.5", "--no-synthetic-tokens"])
    assert b.inner_steps == 500 and b.outer_lr == 0.5 and b.synthetic_tokens is False


def test_loop_bookkeeping_on_cpu(monkeypatch):
    pytest.importorskip("transformers")
    monkeypatch.setattr(torch.cuda, "is_available", lambda: False)
    stats = train_model(diloco_factory=ReferenceDilocoModelPort, **KW)
    assert stats["inner_steps_run"] == 6 and stats["outer_steps_run"] == 3        # every 2 inner steps
    assert len(stats["losses"]) == 12                                             # 2 micro-batches per step
    assert stats["tokens_per_s"] > 0 and np.isfinite(stats["final_loss"])


@pytest.mark.gpu
def test_loop_drop_in_against_reference_port_on_gpu():
    """same loop, same seeds, wrapper swapped; parameters after 3 outer rounds.
    fp32 forward/backward on the GPU is not bit-reproducible between runs of different optimizers'
    memory layouts (atomics in the embedding backward), hence a tolerance."""
    pytest.importorskip("transformers")
    got, ref = {}, {}

    def grab(store, cls):
        def factory(model, *a):
            d = cls(model, *a)
            store["d"] = d
            return d
        return factory

    from nanodiloco_b200 import Diloco
    s1 = train_model(diloco_factory=grab(got, Diloco), **KW)
    s2 = train_model(diloco_factory=grab(ref, ReferenceDilocoModelPort), **KW)
    assert s1["outer_steps_run"] == s2["outer_steps_run"] == 3
    pa = torch.cat([p.detach().reshape(-1) for p in got["d"].model.parameters()])
    pb = torch.cat([p.detach().reshape(-1) for p in ref["d"].model.parameters()])
    torch.testing.assert_close(pa, pb, rtol=1e-4, atol=1e-5)
    assert s1["final_loss"] == pytest.approx(s2["final_loss"], rel=1e-3)
    assert got["d"].avg_sync_time > 0
