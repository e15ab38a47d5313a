"""Checkpoint-file handling of the `clip` face on CPU: TorchScript archives (OpenAI's distribution format), plain
state_dict pickles, fine-tuned `{'model_state_dict': {'clip_model.*', 'classifier.*'}}` checkpoints
(embedding_service.py:16-24, :112-113), geometry inference, refusal of arbitrary pickles."""
import importlib
import os

import pytest
import torch

from conftest import PKG


@pytest.fixture(scope="module")
def clip():
    return importlib.import_module(PKG + ".clip")


def _tiny():
    return {"visual.conv1.weight": torch.randn(8, 3, 2, 2), "visual.transformer.resblocks.0.ln_1.weight": torch.ones(8),
            "logit_scale": torch.tensor(2.5), "context_length": torch.tensor(77)}


def test_script_archive_and_pickles_open_the_same(clip, tmp_path):
    sd = _tiny()
    root = torch.nn.Module()
    for key, value in sd.items():
        parts, m = key.split("."), root
        for p in parts[:-1]:
            if not hasattr(m, p):
                m.add_module(p, torch.nn.Module())
            m = getattr(m, p)
        if value.is_floating_point():
            m.register_parameter(parts[-1], torch.nn.Parameter(value.clone(), requires_grad=False))
        else:
            m.register_buffer(parts[-1], value.clone())
    torch.jit.save(torch.jit.script(root), str(tmp_path / "script.pt"))
    torch.save(sd, str(tmp_path / "plain.pt"))
    torch.save({"epoch": 3, "model_state_dict": {**{"clip_model." + k: v for k, v in sd.items()},
                                                 "classifier.0.weight": torch.zeros(4, 8)},
                "optimizer_state_dict": {"state": {}, "param_groups": [{"lr": 1e-5}]}}, str(tmp_path / "finetuned.pt"))
    for name in ("script.pt", "plain.pt", "finetuned.pt"):
        got = clip._unwrap(clip._load_checkpoint_file(str(tmp_path / name)))
        assert set(got) == set(sd), name
        assert all(torch.equal(got[k], sd[k]) for k in sd), name


def test_arbitrary_pickles_are_refused(clip, tmp_path, monkeypatch):
    import argparse
    torch.save({"state_dict": _tiny(), "extra": argparse.Namespace(lr=1)}, str(tmp_path / "odd.pt"))
    monkeypatch.delenv("EVR_CLIP_TRUSTED_CHECKPOINTS", raising=False)
    with pytest.raises(RuntimeError, match="EVR_CLIP_TRUSTED_CHECKPOINTS"):
        clip._load_checkpoint_file(str(tmp_path / "odd.pt"))


def test_geometry_is_read_off_the_tensors(clip):
    def shapes(width, patch, layers):
        sd = {"visual.conv1.weight": torch.empty((width, 3, patch, patch), device="meta")}
        for i in range(layers):
            sd[f"visual.transformer.resblocks.{i}.attn.in_proj_weight"] = torch.empty((3 * width, width), device="meta")
        return sd
    assert clip.infer_model_name(shapes(768, 32, 12)) == "ViT-B/32"
    assert clip.infer_model_name(shapes(768, 16, 12)) == "ViT-B/16"
    assert clip.infer_model_name(shapes(1024, 14, 24)) == "ViT-L/14"
    with pytest.raises(RuntimeError, match="unsupported vision tower"):
        clip.infer_model_name(shapes(1280, 14, 32))
    with pytest.raises(RuntimeError, match="no vision tower"):
        clip.infer_model_name({"token_embedding.weight": torch.empty(1)})


def test_random_init_is_announced(clip, capsys):
    with pytest.raises(RuntimeError, match="CUDA device only"):
        clip.load("ViT-B/32", device="cpu", random_init=3)
    assert "RANDOM weights" in capsys.readouterr().err
