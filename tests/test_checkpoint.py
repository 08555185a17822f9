"""CPU: checkpoint / config compatibility.  A state_dict saved by the REFERENCE module (staged in
oracle/_ref by oracle/build_ref.py) with `torch.save` loads into this package's Decoder through the reference's own
`load_model` protocol (generate_caption.py:38-69): strict first, `strict=False` fallback with the printed notice; and a
checkpoint written by this package loads strictly into the reference module."""
import argparse
import json
import os

import pytest
import torch

from oracle import ref_runner
from sat_b200 import checkpoint as CK

needs_ref = pytest.mark.skipif(not ref_runner.available(), reason="oracle/_ref not staged (run python oracle/build_ref.py)")


def _config(tmp_path, **over):
    wd = {"<start>": 0, "<eos>": 1, "<unk>": 2, "<pad>": 3}
    wd.update({f"w{i}": 4 + i for i in range(96)})
    data = tmp_path / "data"
    data.mkdir(exist_ok=True)
    (data / "word_dict.json").write_text(json.dumps(wd))
    ns = argparse.Namespace(batch_size=64, epochs=10, lr=1e-4, step_size=5, alpha_c=1, perform_test=True, seed=42, log_interval=100,
                            data=str(data), network="vgg19", model=None, tf=True, ado=True, fraction=1.0, bert=False, attention=True)
    for k, v in over.items():
        setattr(ns, k, v)
    cfg = tmp_path / "model" / "model_config.json"
    CK.save_model_config(str(cfg), ns)                         # train.py:108-109
    return ns, str(cfg), len(wd)


@needs_ref
def test_reference_checkpoint_round_trip(tmp_path):
    ns, cfg, V = _config(tmp_path)
    torch.manual_seed(0)
    ref = ref_runner.ref_decoder_module("cpu").Decoder(V, 512, tf=True, ado=True, bert=False, attention=True)
    path = str(tmp_path / "model" / "model_vgg19_3.pth")
    torch.save(ref.state_dict(), path)                          # train.py:103-104
    dec, bert, mp, cp = CK.load_model(path, cfg)
    assert not bert and mp == path and cp == cfg and not dec.training
    assert dec.use_advanced_deep_output and dec.use_attention and not dec.use_tf and dec.vocabulary_size == V
    sd_ref, sd = ref.state_dict(), dec.state_dict()
    assert list(sd_ref.keys()) == list(sd.keys())
    for k in sd_ref:
        assert torch.equal(sd_ref[k], sd[k]), k
    # and back: our checkpoint loads strictly into the reference module
    out = str(tmp_path / "model" / "ours.pth")
    CK.save_checkpoint(dec, out)
    ref2 = ref_runner.ref_decoder_module("cpu").Decoder(V, 512, ado=True, attention=True)
    ref2.load_state_dict(torch.load(out))
    for k, v in ref2.state_dict().items():
        assert torch.equal(v, sd_ref[k]), k


@needs_ref
def test_strict_fallback_like_generate_caption(tmp_path, capsys):
    # a checkpoint trained WITHOUT --ado, loaded with a config that says ado: f_h / f_z / f_out keys are missing
    ns, cfg, V = _config(tmp_path, ado=True)
    ref = ref_runner.ref_decoder_module("cpu").Decoder(V, 512, tf=True, ado=False, attention=True)
    path = str(tmp_path / "model" / "noado.pth")
    torch.save(ref.state_dict(), path)
    dec, *_ = CK.load_model(path, cfg)
    assert "Strict loading failed, loading with strict=False" in capsys.readouterr().out      # generate_caption.py:68
    assert torch.equal(dec.lstm.weight_ih, ref.lstm.weight_ih) and hasattr(dec, "f_out")


def test_load_model_argument_errors(tmp_path):
    with pytest.raises(ValueError):
        CK.load_model(None, None)
    ns, cfg, V = _config(tmp_path, network="alexnet")
    torch.save({}, str(tmp_path / "m.pth"))
    with pytest.raises(ValueError):
        CK.load_model(str(tmp_path / "m.pth"), cfg)
    # explicit vocabulary size when <data>/word_dict.json is not around
    ns, cfg, V = _config(tmp_path, data=str(tmp_path / "nowhere"), ado=False)
    from sat_b200.decoder import Decoder
    d = Decoder(77, 512, attention=True)
    CK.save_checkpoint(d, str(tmp_path / "m2.pth"))
    dec, *_ = CK.load_model(str(tmp_path / "m2.pth"), cfg, vocabulary_size=77)
    assert dec.vocabulary_size == 77 and torch.equal(dec.embedding.weight, d.embedding.weight)
    assert os.path.exists(cfg)
