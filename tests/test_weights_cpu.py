"""Checkpoint loading: a file with the reference's state-dict keys
(concept_attention/image_generator.py:19-48) loads into the engine's dict; wrong files fail loudly."""
import pytest
import torch

from conceptattention_b200.weights import CONFIGS, expected_keys, init_state_dict, load_checkpoint


def test_safetensors_and_pt_roundtrip(tmp_path):
    from safetensors.torch import save_file
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=1, dtype=torch.bfloat16)
    assert set(sd) == set(expected_keys(p))
    assert all(tuple(sd[k].shape) == s for k, s in expected_keys(p).items())
    f1, f2 = tmp_path / "tiny.safetensors", tmp_path / "tiny.pt"
    save_file(sd, str(f1))
    torch.save(sd, str(f2))
    for f in (f1, f2):
        got = load_checkpoint(f, p)
        assert set(got) == set(sd) and all(torch.equal(got[k], sd[k]) for k in sd)


def test_mismatched_checkpoints_are_rejected(tmp_path):
    from safetensors.torch import save_file
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=1, dtype=torch.bfloat16)
    missing = {k: v for k, v in sd.items() if k != "img_in.weight"}
    save_file(missing, str(tmp_path / "missing.safetensors"))
    with pytest.raises(ValueError, match="missing"):
        load_checkpoint(tmp_path / "missing.safetensors", p)
    extra = dict(sd, **{"pe_embedder.buffer": torch.zeros(2)})
    save_file(extra, str(tmp_path / "extra.safetensors"))
    with pytest.raises(ValueError, match="unexpected"):
        load_checkpoint(tmp_path / "extra.safetensors", p)
    assert set(load_checkpoint(tmp_path / "extra.safetensors", p, strict=False)) == set(sd)
    with pytest.raises(ValueError, match="wrong shape"):                # dev checkpoint into schnell params etc.
        load_checkpoint(tmp_path / "extra.safetensors", CONFIGS["tiny-dev"], strict=False)
