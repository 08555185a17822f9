"""Checkpoint and config compatibility with the reference.

The reference saves `torch.save(decoder.state_dict(), f'model/model_{network}_{epoch}.pth')` plus
`model/model_config.json = json.dump(vars(args))` after every epoch (train.py:103-110) and reloads them in
generate_caption.py:24-74 (`load_model`): read `network / data / ado / bert / attention` from the config, size the
vocabulary from `<data>/word_dict.json` (or BERT's), build `Decoder(vocabulary_size, encoder.dim, ado=, bert=, attention=)`
and `load_state_dict` strict, falling back to `strict=False` with a printed notice.  The module tree of this package's
`Decoder` has the reference's keys and shapes, so those files load as they are; this module restates the two small
host-side routines around them.  The CNN encoder is out of scope: only its feature width is needed (encoder.py:16,21,26).
"""
from __future__ import annotations

import json
import os

import torch

# encoder.py:13-26: feature width per --network choice
ENCODER_DIMS = {"vgg19": 512, "resnet152": 2048, "densenet161": 1920}


def save_checkpoint(decoder, model_path: str):
    """train.py:103-104: the decoder's state_dict, reference key names, fp32 tensors."""
    d = os.path.dirname(model_path)
    if d:
        os.makedirs(d, exist_ok=True)
    torch.save({k: v.detach().cpu().clone() for k, v in decoder.state_dict().items()}, model_path)


def save_model_config(config_path: str, args):
    """train.py:108-109: `json.dump(vars(args))` (accepts an argparse Namespace or a dict)."""
    d = os.path.dirname(config_path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(config_path, "w") as f:
        json.dump(dict(args) if isinstance(args, dict) else vars(args), f)


def load_model(model_path=None, model_config_path=None, *, vocabulary_size=None, compute_dtype=torch.float32,
               device=None, gate=True):
    """generate_caption.py:24-74 for the decoder (the path in scope).  Returns (decoder, bert, model_path,
    model_config_path) with the decoder in eval mode (generate_caption.py:71-72).  `vocabulary_size` overrides the
    `<data>/word_dict.json` lookup (generate_caption.py:59-60) for callers that hold the vocabulary elsewhere."""
    from .decoder import Decoder
    if model_path is None or model_config_path is None:
        raise ValueError("Model path and config path must be provided if not loading from wandb")     # generate_caption.py:35-36
    with open(model_config_path, "r") as f:
        model_config = json.load(f)
    network = model_config["network"]
    ado, bert, attention = model_config["ado"], model_config["bert"], model_config["attention"]
    if network not in ENCODER_DIMS:
        raise ValueError(f"unknown encoder network '{network}'")
    if bert:
        vocabulary_size = vocabulary_size or 0          # Decoder(bert=True) takes the size from the BERT config (decoder.py:26)
    elif vocabulary_size is None:
        with open(os.path.join(model_config["data"], "word_dict.json"), "r") as f:
            vocabulary_size = len(json.load(f))
    decoder = Decoder(vocabulary_size, ENCODER_DIMS[network], ado=ado, bert=bert, attention=attention,
                      compute_dtype=compute_dtype, gate=gate)
    state = torch.load(model_path, map_location="cpu")
    try:
        decoder.load_state_dict(state)
    except RuntimeError:
        print("Strict loading failed, loading with strict=False")                                      # generate_caption.py:68
        decoder.load_state_dict(state, strict=False)
    if device is not None:
        decoder = decoder.to(device)
    decoder.eval()
    return decoder, bert, model_path, model_config_path
