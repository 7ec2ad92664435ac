"""Parameter layout of the fused GPT-2 engine (no GPU needed: only the planning code runs).

Gradient segments must follow the order in which backward completes them, because each segment is reduced and stepped as
soon as its last contributor has been enqueued: a tensor placed in too early a segment would be reduced before its
gradient is complete."""
import math

import pytest

from trustworthy_dl.engine.gpt2_engine import SEG_ALIGN, GPT2Engine
from trustworthy_dl.models.gpt2 import GPT2Config


def _plan(n_layer, layers, has_embed, has_head, tied, n_buckets):
    cfg = GPT2Config(vocab_size=1000, padded_vocab_size=1024, n_positions=128, n_embd=128, n_layer=n_layer, n_head=2)
    e = GPT2Engine.__new__(GPT2Engine)                     # planning only: nothing is allocated
    e.cfg, e.C, e.V, e.Vp = cfg, cfg.n_embd, cfg.vocab_size, cfg.padded_vocab_size
    e.l0, e.l1 = layers
    e.L = e.l1 - e.l0
    e.has_embed, e.has_head = has_embed, has_head
    e.tied_head = tied and has_embed and has_head
    e.n_buckets = max(1, min(n_buckets, e.L))
    return e, e._plan_segments()


@pytest.mark.parametrize("n_layer,layers,has_embed,has_head,tied,n_buckets", [
    (24, (0, 24), True, True, True, 4), (24, (0, 24), True, True, False, 4), (24, (0, 6), True, False, False, 2),
    (24, (6, 13), False, False, False, 2), (24, (20, 24), False, True, False, 1), (2, (0, 2), True, True, True, 4),
    (5, (0, 5), True, True, True, 3)])
def test_segments_follow_gradient_completion_order(n_layer, layers, has_embed, has_head, tied, n_buckets):
    e, segs = _plan(n_layer, layers, has_embed, has_head, tied, n_buckets)
    seg_of = {name: i for i, seg in enumerate(segs) for name, _, _ in seg.tensors}
    # every tensor exactly once
    names = [name for seg in segs for name, _, _ in seg.tensors]
    assert len(names) == len(set(names))
    want = {f"h{l}.{k}" for l in range(*layers) for k in ("qkv_w", "o_w", "fc_w", "proj_w", "ln1_w", "ln1_b", "qkv_b", "o_b",
                                                            "ln2_w", "ln2_b", "fc_b", "proj_b")}
    if has_embed:
        want |= {"wte", "wpe"}
    if has_head:
        want |= {"lnf_w", "lnf_b"} | (set() if e.tied_head else {"lm_head"})
    assert set(names) == want
    # backward walks the blocks from the top down: a higher block never sits in a later segment than a lower one,
    # and the segment index handed to `on_segment_done` is the one the block's tensors live in
    for l in range(layers[0], layers[1]):
        assert seg_of[f"h{l}.qkv_w"] == e._seg_of_layer[l]
        assert all(seg_of[f"h{l}.{k}"] == e._seg_of_layer[l] for k in ("proj_b", "ln1_w", "fc_b", "o_w"))
        if l + 1 < layers[1]:
            assert e._seg_of_layer[l + 1] <= e._seg_of_layer[l]
    # ln_f / an untied head finish first (they ride with the top block group); the embeddings finish last -- the tied
    # wte receives its final contribution from the embedding backward
    if has_head:
        assert seg_of["lnf_w"] == 0 and (e.tied_head or seg_of["lm_head"] == 0)
    if has_embed:
        assert seg_of["wte"] == seg_of["wpe"] == len(segs) - 1 and all(seg_of[n] < len(segs) - 1 for n in names if n not in ("wte", "wpe"))
    # decayed tensors first inside a segment (AdamW takes one boundary per call)
    for seg in segs:
        flags = [d for _, _, d in seg.tensors]
        assert flags == sorted(flags, reverse=True)
        assert all(math.prod(s) % 8 == 0 for _, s, _ in seg.tensors)
    assert SEG_ALIGN % (8 * 8) == 0                      # 16-byte vectors x up to 8 ranks divide every segment
