import pytest
import torch

from trustworthy_dl import ModelFactory, get_model
from trustworthy_dl.core.partition import StageMap, balanced_partition, build_stages
from trustworthy_dl.utils.data_loader import get_dataloader


def test_factory_names_and_param_counts():
    f = ModelFactory()
    assert {"gpt2", "gpt2-medium", "resnet32", "resnet50", "resnet101", "vgg11", "vgg13", "vgg16"} <= set(f.available_models())
    r32 = f.create_model("resnet32")
    assert 0.45e6 < r32.num_parameters() < 0.48e6
    with pytest.raises(ValueError):
        f.create_model("alexnet")


@pytest.mark.parametrize("name,shape,classes", [("resnet32", (2, 3, 32, 32), 10), ("vgg11-cifar", (2, 3, 32, 32), 10)])
def test_vision_forward_and_stage_equivalence(name, shape, classes):
    torch.manual_seed(0)
    m = get_model(name).eval()
    x = torch.randn(*shape)
    y = m(x)
    assert y.shape == (2, classes)
    for k in (2, 3, 8):
        z = x
        for st in build_stages(m, k):
            z = st(z)
        assert torch.allclose(z, y, atol=1e-5)


def test_gpt2_units_cover_everything_and_chain():
    torch.manual_seed(0)
    m = get_model("gpt2-tiny").eval()
    assert hasattr(m.transformer, "h") and len(m.transformer.h) == 4      # reference contract (:126)
    idx = torch.randint(0, 512, (2, 16))
    y = m(idx)
    assert y.shape == (2, 16, 512)
    stages = build_stages(m, 3)
    z = idx
    for st in stages:
        z = st(z)
    assert torch.allclose(z, y, atol=1e-5)
    covered = {id(p) for st in stages for p in st.parameters()}
    assert covered == {id(p) for p in m.parameters()}                      # Q3: nothing left on no stage


def test_gpt2_medium_shape_contract():
    from trustworthy_dl.models.gpt2 import GPT2Config, GPT2_PRESETS
    cfg = GPT2Config(**GPT2_PRESETS["gpt2-medium"])
    assert (cfg.n_embd, cfg.n_layer, cfg.n_head, cfg.padded_vocab_size) == (1024, 24, 16, 50304)
    per_block = 12 * 1024 * 1024 + 13 * 1024
    assert per_block == 12_596_224


def test_balanced_partition_properties():
    assert balanced_partition([1, 1, 1, 1], 2) == [0, 2, 4]
    assert balanced_partition([10, 1, 1, 1, 1], 2) == [0, 1, 5]
    b = balanced_partition([3, 1, 4, 1, 5, 9, 2, 6], 8)
    assert b == list(range(9))
    b = balanced_partition([1.0] * 26, 4)                                   # remainder not dropped
    sizes = [j - i for i, j in zip(b[:-1], b[1:])]
    assert sum(sizes) == 26 and max(sizes) - min(sizes) <= 1


def test_stage_map_moves():
    sm = StageMap(4, 4)
    assert sm.owner == [0, 1, 2, 3]
    assert sm.move_all(1, 3) == [1] and sm.stages_of(3) == [1, 3] and sm.stages_of(1) == []


def test_synthetic_loaders():
    dl = get_dataloader("openwebtext", "train", 2, num_batches=2, seq_len=32)
    b = next(iter(dl))
    assert b["input"].shape == (2, 32) and b["target"].shape == (2, 32) and b["input"].dtype == torch.int64
    assert torch.equal(b["input"][:, 1:], b["target"][:, :-1])
    assert int(b["input"].max()) < 50257
    b2 = next(iter(get_dataloader("openwebtext", "train", 2, num_batches=2, seq_len=32)))
    assert torch.equal(b["input"], b2["input"])                            # deterministic
    c = next(iter(get_dataloader("cifar10", "validation", 4)))
    assert c["input"].shape == (4, 3, 32, 32) and c["target"].shape == (4,)
    i = next(iter(get_dataloader("imagenet", "train", 1, num_batches=1)))
    assert i["input"].shape == (1, 3, 224, 224)
