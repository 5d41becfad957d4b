"""``build_model(args)`` with the reference's signature and return value (builder.py:49-56):

    model, tokenizer, point_backbone_config, mm_use_point_start_end = build_model(args)

``args`` is the argparse namespace of train.py:313-347 / evaluate.py:175-188; the fields read here are ``model_name``,
``num_bins``, ``unfreeze_pc_encoder``, ``unfreeze_language_model`` (+ optional ``device``).  When ``args.model_name`` is a
local HF checkpoint directory the config, tokenizer and weights come from it exactly like init_model (builder.py:9-30);
offline (no checkpoint: the PointLLM weights are a hub download, model_arch.py:29) ``model_name`` may be ``"synthetic:full"``,
``"synthetic:tiny"`` or ``"synthetic:full_base"`` - the named architecture with random-init weights and a minimal tokenizer
that knows the special / trajectory tokens (enough for collate_fn's pad id and the ``<pK>`` decode of train.py:240-246).
"""
from __future__ import annotations

import os
import re
from typing import List, Sequence

import torch

from .config import TrajVLMConfig, POINT_PATCH_TOKEN_ID, POINT_START_TOKEN_ID, POINT_END_TOKEN_ID
from .model import TrajPointLLMForCausalLM

# constant.py of the reference (token templates)
RT2_TOKEN_TEMPLATE = "<p{p}>"
TIMESTEP_START_TOKEN, TIMESTEP_SEP_TOKEN, TIMESTEP_END_TOKEN = "<ts>", "<tsep>", "<te>"
BASE_VOCAB = 32000


class SyntheticTokenizer:
    """Stand-in for the LLaMA tokenizer when no checkpoint directory is available: ids 0..31999 are opaque words, the
    point / trajectory specials sit where builder.py:33-46 puts them.  Implements what the entry points touch:
    ``pad_token_id`` / ``eos_token_id`` / ``bos_token_id``, ``len()``, ``add_tokens``, ``convert_tokens_to_ids``,
    ``decode`` / ``batch_decode`` (``<pK>`` tokens separated by a blank like sentencepiece pieces, train.py:240-246)."""

    pad_token_id, bos_token_id, eos_token_id = 0, 1, 2

    def __init__(self):
        self.special = {"<point_patch>": POINT_PATCH_TOKEN_ID, "<point_start>": POINT_START_TOKEN_ID, "<point_end>": POINT_END_TOKEN_ID}
        self._n = BASE_VOCAB + 3

    def __len__(self):
        return self._n

    def add_tokens(self, tokens: Sequence[str]) -> int:
        for t in tokens:
            if t not in self.special:
                self.special[t] = self._n
                self._n += 1
        return len(tokens)

    def convert_tokens_to_ids(self, tokens):
        if isinstance(tokens, str):
            return self.special.get(tokens, 0)
        return [self.special.get(t, 0) for t in tokens]

    def decode(self, ids, skip_special_tokens: bool = False) -> str:
        inv = {v: k for k, v in self.special.items()}
        out, prev_bin = [], False
        for i in (ids.tolist() if torch.is_tensor(ids) else list(ids)):
            i = int(i)
            if i in (self.pad_token_id, self.bos_token_id, self.eos_token_id):
                if not skip_special_tokens:
                    out.append({0: "<unk>", 1: "<s>", 2: "</s>"}[i])
                prev_bin = False
                continue
            s = inv.get(i, f"w{i}")
            is_bin = bool(re.fullmatch(r"<p\d+>", s))
            out.append((" " if (prev_bin and is_bin) else "") + s)
            prev_bin = is_bin
        return "".join(out)

    def batch_decode(self, seqs, skip_special_tokens: bool = False) -> List[str]:
        return [self.decode(s, skip_special_tokens) for s in seqs]


def _synthetic_config(kind: str, vocab: int) -> TrajVLMConfig:
    cfg = {"full": TrajVLMConfig.full, "tiny": TrajVLMConfig.tiny, "full_base": TrajVLMConfig.full_base,
           "tiny_base": TrajVLMConfig.tiny_base, "tiny_hd128": TrajVLMConfig.tiny_hd128}[kind]()
    cfg.vocab_size = vocab
    return cfg


def init_model(args):
    """builder.py:9-30."""
    model_name = os.path.expanduser(str(args.model_name))
    num_bins = int(getattr(args, "num_bins", 256))
    vocab = BASE_VOCAB + 3 + (3 + num_bins if num_bins > 0 else 0)           # builder.py:37-44
    device = getattr(args, "device", "cuda")
    if model_name.startswith("synthetic:"):
        config = _synthetic_config(model_name.split(":", 1)[1], vocab)
        tokenizer = SyntheticTokenizer()
        model = TrajPointLLMForCausalLM(args, config, None, device=device)
        g = torch.Generator(device=model.dev).manual_seed(int(getattr(args, "seed", 0) or 0))
        with torch.no_grad():                                                 # HF initializer_range for the LLM, fan-in for the point side
            for name, p in list(model.named_parameters()) + list(model.named_buffers()):
                leaf = name.rsplit(".", 1)[-1]
                if leaf == "num_batches_tracked":
                    continue
                if leaf == "running_var" or ("norm" in name and leaf == "weight") or (".1.weight" in name and "conv" in name):
                    p.fill_(1.0)
                elif leaf in ("running_mean", "bias"):
                    p.zero_()
                elif p.dim() >= 2 and ("point_backbone" in name or "point_proj" in name):
                    p.normal_(0.0, 1.0 / (p.shape[1] ** 0.5), generator=g)
                else:
                    p.normal_(0.0, 0.02, generator=g)
        model.repack()
    else:
        from transformers import AutoConfig, AutoTokenizer
        hf = AutoConfig.from_pretrained(model_name)
        hf.vocab_size = vocab                                                 # allocate at the post-add_trajectory_token size
        tokenizer = AutoTokenizer.from_pretrained(model_name)
        model = TrajPointLLMForCausalLM(args, hf, model_name, device=device)
    model.initialize_tokenizer_point_backbone_config_wo_embedding(tokenizer)
    mm_use_point_start_end = bool(getattr(model.cfg, "mm_use_point_start_end", False))
    return model, tokenizer, model.get_model().point_backbone_config, mm_use_point_start_end


def add_trajectory_token(args, model, tokenizer):
    """builder.py:33-46."""
    if int(getattr(args, "num_bins", 256)) > 0:
        tokenizer.add_tokens([TIMESTEP_START_TOKEN, TIMESTEP_SEP_TOKEN, TIMESTEP_END_TOKEN])
        tokenizer.add_tokens([RT2_TOKEN_TEMPLATE.format(p=p) for p in range(int(args.num_bins))])
    model.resize_token_embeddings(len(tokenizer), mean_resizing=False)
    return model, tokenizer


def build_model(args):
    """builder.py:49-56."""
    model, tokenizer, point_backbone_config, mm_use_point_start_end = init_model(args)
    model, tokenizer = add_trajectory_token(args, model, tokenizer)
    return model, tokenizer, point_backbone_config, mm_use_point_start_end
