"""Drop-in module for the reference's ``TrajPointLLMForCausalLM`` (egoscaler/models/pointllm/model_arch.py:8-124),
i.e. ``PointLLMLlamaForCausalLM`` (pointllm/model/pointllm.py:181-275) with EgoScaler's freezing semantics, running on
the hand-written CUDA path (no ATen math, no CPU fallback).

Same surface: ``forward(input_ids, attention_mask, ..., point_clouds, return_dict)`` -> object with ``.logits``;
``generate(...)`` -> object with ``.sequences`` / ``.scores``; ``train(mode)`` keeps frozen parts in eval;
``state_dict()`` / ``load_state_dict()`` use the reference's parameter names (params.ParamTree);
``get_model().point_backbone_config`` carries the same keys (pointllm.py:49-59,281-300).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from . import ops
from .config import (TrajVLMConfig, POINT_PATCH_TOKEN_ID, POINT_START_TOKEN_ID, POINT_END_TOKEN_ID, config_from_hf)
from .encoder import (PackedPointEncoder, PackedProjector, point_encoder_forward, projector_forward)
from .llm import KVCache, LayerWeights, LlamaTrunk
from .params import ParamTree


@dataclass
class CausalLMOutput:
    """Stand-in for transformers' CausalLMOutputWithPast (only the fields the reference reads)."""
    logits: torch.Tensor
    past_key_values: Optional[KVCache] = None
    loss: Optional[torch.Tensor] = None


@dataclass
class GenerateOutput:
    """Stand-in for GenerateDecoderOnlyOutput: ``.sequences`` [B, P+T], ``.scores`` tuple of T x [B, V] fp32."""
    sequences: torch.Tensor
    scores: Optional[Tuple[torch.Tensor, ...]] = None


class _ModelView:
    """What ``get_model()`` returns in the reference: the inner model carrying ``point_backbone_config``."""

    def __init__(self, outer: "TrajPointLLMForCausalLM"):
        self._outer = outer
        self.point_backbone_config = outer.point_backbone_config


class TrajPointLLMForCausalLM(ParamTree):
    def __init__(self, *args, **kwargs):
        """Two call forms:

        * the reference's ``TrajPointLLMForCausalLM(args, config, model_name)`` (model_arch.py:13-24): ``args`` carries
          ``unfreeze_pc_encoder`` / ``unfreeze_language_model`` (train.py:344-345), ``config`` is the HF / PointLLM config
          object (``config_from_hf``), ``model_name`` a checkpoint directory or file whose ``state_dict`` is loaded when it
          exists (the reference's ``from_pretrained`` at model_arch.py:26-31); optional ``device=``;
        * the native ``TrajPointLLMForCausalLM(cfg: TrajVLMConfig, device="cuda", unfreeze_pc_encoder=False,
          unfreeze_language_model=False)``.
        """
        ref_args = model_name = hf_config = None
        if args and isinstance(args[0], TrajVLMConfig):
            cfg = args[0]
            device = args[1] if len(args) > 1 else kwargs.pop("device", "cuda")
            unfreeze_pc_encoder = bool(kwargs.pop("unfreeze_pc_encoder", False))
            unfreeze_language_model = bool(kwargs.pop("unfreeze_language_model", False))
        else:
            names = ("args", "config", "model_name")
            vals = dict(zip(names, args))
            vals.update({k: kwargs.pop(k) for k in names if k in kwargs})
            ref_args, hf_config, model_name = vals.get("args"), vals.get("config"), vals.get("model_name")
            cfg = hf_config if isinstance(hf_config, TrajVLMConfig) else config_from_hf(hf_config)
            device = kwargs.pop("device", "cuda")
            unfreeze_pc_encoder = bool(getattr(ref_args, "unfreeze_pc_encoder", False))
            unfreeze_language_model = bool(getattr(ref_args, "unfreeze_language_model", False))
        if kwargs:
            raise TypeError(f"unexpected arguments: {sorted(kwargs)}")
        super().__init__(cfg, device)
        self.args, self.config, self.model_name = ref_args, hf_config if hf_config is not None else cfg, model_name
        self.cfg = cfg
        self.dev = torch.device(device)
        p = cfg.point
        # pointllm.py:49-59 + 281-300 (ids are fixed by the vocabulary layout of builder.py:33-46)
        self.point_backbone_config = {
            "point_cloud_dim": p.point_dims, "backbone_output_dim": p.trans_dim * (2 if p.use_max_pool else 1),
            "project_output_dim": cfg.hidden_size, "point_token_len": p.point_token_len,
            "mm_use_point_start_end": cfg.mm_use_point_start_end, "projection_hidden_layer": len(p.projection_hidden_dim),
            "use_max_pool": p.use_max_pool, "default_point_patch_token": "<point_patch>",
            "default_point_start_token": "<point_start>", "default_point_end_token": "<point_end>",
            "point_patch_token": POINT_PATCH_TOKEN_ID, "point_start_token": POINT_START_TOKEN_ID,
            "point_end_token": POINT_END_TOKEN_ID,
        }
        self.unfreeze_pc_encoder = unfreeze_pc_encoder
        self.unfreeze_language_model = unfreeze_language_model
        self._configure_trainable_parameters()
        self._packed_encoder: Optional[PackedPointEncoder] = None
        self._packed_proj: Optional[PackedProjector] = None
        self._trunk: Optional[LlamaTrunk] = None
        self._lm_head_T: Optional[torch.Tensor] = None
        self._lm_head_tiled = None                          # decode-layout copy of lm_head (ops.retile_weight)
        self._lm_head_folded = False                        # ... with model.norm folded into its columns
        self.v_pad = (cfg.vocab_size + 7) // 8 * 8          # logits leading dimension (16-byte rows)
        self._enc_train = None                              # train-mode point encoder operands (--unfreeze_pc_encoder)
        self._versions = None                               # autograd version counters the derived operands were built from
        if model_name is not None:
            self.load_pretrained_weights()

    # ------------------------------------------------------------------ checkpoint (model_arch.py:26-31)
    def load_pretrained_weights(self):
        """``load_state_dict(parent_state_dict, strict=False)`` of model_arch.py:26-31 from ``model_name``: a directory
        holding ``pytorch_model*.bin`` / ``*.safetensors`` shards or a single state-dict file.  A name that is not a path
        (a hub id: no network here) leaves the random initialisation in place and says so."""
        import glob
        import os
        import warnings
        path = os.path.expanduser(str(self.model_name))
        files = []
        if os.path.isdir(path):
            files = sorted(glob.glob(os.path.join(path, "*.safetensors"))) or sorted(glob.glob(os.path.join(path, "pytorch_model*.bin")))
        elif os.path.isfile(path):
            files = [path]
        if not files:
            warnings.warn(f"egoscaler_b200: no checkpoint files under {path!r}; parameters keep their initial values")
            return
        own = self.state_dict()
        for f in files:
            if f.endswith(".safetensors"):
                from safetensors.torch import load_file
                part = load_file(f)
            else:
                part = torch.load(f, map_location="cpu", weights_only=True)
            for k, v in part.items():
                if k not in own:
                    continue
                t = own[k]
                if t.dim() == 2 and v.dim() == 2 and v.shape[0] < t.shape[0] and v.shape[1] == t.shape[1]:
                    t[:v.shape[0]].copy_(v)                 # vocabulary grown by add_trajectory_token (builder.py:33-46)
                else:
                    t.copy_(v.view(t.shape))
        self.repack()

    def resize_token_embeddings(self, new_num_tokens: int, mean_resizing: bool = False):
        """builder.py:44.  The parameter tree is allocated at ``cfg.vocab_size`` rows already (32262 after
        add_trajectory_token); rows a smaller checkpoint did not provide keep their N(0, 0.02) initialisation, which is
        what HF does for ``mean_resizing=False``.  Any other size would need a re-allocation: refuse loudly."""
        if int(new_num_tokens) != self.cfg.vocab_size:
            raise ValueError(f"egoscaler_b200 allocates embed_tokens / lm_head at cfg.vocab_size={self.cfg.vocab_size}; "
                             f"build the config with vocab_size={int(new_num_tokens)} instead of resizing")
        return self.model.embed_tokens

    def initialize_tokenizer_point_backbone_config_wo_embedding(self, tokenizer):
        """pointllm.py:281-300: read the point-token ids from the tokenizer into ``point_backbone_config``."""
        c = self.point_backbone_config
        for key, tok in (("point_patch_token", c["default_point_patch_token"]), ("point_start_token", c["default_point_start_token"]),
                         ("point_end_token", c["default_point_end_token"])):
            if hasattr(tokenizer, "convert_tokens_to_ids"):
                c[key] = int(tokenizer.convert_tokens_to_ids([tok])[0])

    def enc_train(self):
        """Operands of the train-mode point encoder (BatchNorm batch statistics + DropPath), or None when it is frozen /
        the module is in eval mode (model_arch.py:110-124)."""
        if not self.unfreeze_pc_encoder:
            return None
        if self._enc_train is None:
            from .encoder_train import TrainPointEncoder
            self._enc_train = {"pk": TrainPointEncoder(self.cfg, self.named())}
        return self._enc_train

    # ------------------------------------------------------------------ reference API surface
    def get_model(self) -> _ModelView:
        return _ModelView(self)

    def _configure_trainable_parameters(self):
        """model_arch.py:33-51: everything trainable except LLM layers / point encoder unless un-frozen."""
        for n, q in self.named_parameters():
            if n.startswith("model.point_backbone."):
                q.requires_grad = bool(self.unfreeze_pc_encoder)
            elif n.startswith("model.layers."):
                q.requires_grad = bool(self.unfreeze_language_model)
            else:
                q.requires_grad = True

    def train(self, mode: bool = True):
        """model_arch.py:110-124: frozen sub-modules stay in eval.  The CUDA path has no dropout and uses BatchNorm
        running statistics only, which is exactly the reference behaviour while the encoder is frozen.  With
        --unfreeze_pc_encoder the trainer routes the encoder through encoder_train.py (batch statistics + DropPath)."""
        super().train(mode)
        return self

    def trainable_names(self) -> List[str]:
        return [n for n, q in self.named_parameters() if q.requires_grad]

    # ------------------------------------------------------------------ operand packing
    def repack_encoder(self):
        """Rebuild only the point-encoder operands (after an optimizer step on its parameters)."""
        self._packed_encoder = PackedPointEncoder(self.cfg, self.named())

    def repack(self):
        """(Re)build kernel operands from the parameters: call after load_state_dict / optimizer steps on the
        projector / lm_head.  Trunk weights are the parameters themselves (views), nothing is copied."""
        named = self.named()
        self._packed_encoder = PackedPointEncoder(self.cfg, named)
        pw = [named[w] for w, _ in self.projector_params()]
        pb = [named[b] for _, b in self.projector_params()]
        self._packed_proj = PackedProjector(pw, pb)
        layers = []
        for i in range(self.cfg.num_hidden_layers):
            pre = f"model.layers.{i}."
            layers.append(LayerWeights(self.wqkv[i], named[pre + "self_attn.o_proj.weight"].detach(), self.wgu[i],
                                       named[pre + "mlp.down_proj.weight"].detach(),
                                       named[pre + "input_layernorm.weight"].detach(),
                                       named[pre + "post_attention_layernorm.weight"].detach()))
        self._trunk = LlamaTrunk(self.cfg, layers, self.dev)
        self._lm_head_T = None
        self._lm_head_tiled = None
        if self._enc_train is not None:
            self._enc_train["pk"].refresh()
        self._versions = self._version_signature()

    def _version_signature(self):
        named = self.named()
        proj = tuple(named[n]._version for pair in self.projector_params() for n in pair)
        head = (self.lm_head.weight._version,)
        enc = tuple(t._version for n, t in named.items() if n.startswith("model.point_backbone."))
        trunk = tuple((self.wqkv[i]._version, self.wgu[i]._version,
                       named[f"model.layers.{i}.self_attn.o_proj.weight"]._version, named[f"model.layers.{i}.mlp.down_proj.weight"]._version)
                      for i in range(self.cfg.num_hidden_layers))
        return {"proj": proj, "head": head, "enc": enc, "trunk": trunk}

    def sync_derived(self):
        """Rebuild the operand copies derived from parameters that an external optimizer (torch.optim.AdamW / DeepSpeed
        stepping ``model.parameters()``) has updated in place since they were built: fp32 projector biases and transposed
        projector weights, the contraction-major lm_head / trunk copies, the decode-layout copies, the folded point-encoder
        operands.  Detection = autograd version counters (in-place updates bump them); ``Trainer`` refreshes explicitly."""
        if self._trunk is None:
            self.repack()
            return
        sig = self._version_signature()
        old = self._versions
        if old == sig:
            return
        if old is None or old["proj"] != sig["proj"]:
            self._packed_proj.refresh()
        if old is None or old["head"] != sig["head"]:
            self._lm_head_T = None
            self._lm_head_tiled = None
        if old is None or old["enc"] != sig["enc"]:
            self.repack_encoder()
            if self._enc_train is not None:
                self._enc_train["pk"].refresh()
        if old is not None and old["trunk"] != sig["trunk"]:
            for i, lw in enumerate(self._trunk.layers):
                if old["trunk"][i] != sig["trunk"][i]:
                    lw.refresh_derived()
        self._versions = sig

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        r = super().load_state_dict(state_dict, strict=strict, assign=False)
        self.repack()
        return r

    @property
    def trunk(self) -> LlamaTrunk:
        if self._trunk is None:
            self.repack()
        return self._trunk

    def lm_head_T(self) -> torch.Tensor:
        """[d, V] copy of lm_head for the data-gradient GEMM (refreshed by the trainer after each step)."""
        if self._lm_head_T is None:
            self._lm_head_T = ops.transpose(self.lm_head.weight.detach())
        return self._lm_head_T

    # ------------------------------------------------------------------ pieces of the forward
    @staticmethod
    def draw_fps_start(n_points: int, batch: int) -> torch.Tensor:
        """The draw pointbert/misc.py:52 makes on the CPU generator (same stream consumption as the reference)."""
        return torch.randint(0, n_points, (batch,), dtype=torch.long)

    def point_features(self, point_clouds: torch.Tensor, fps_start: Optional[torch.Tensor] = None,
                       save: Optional[list] = None, enc_train=None) -> torch.Tensor:
        """pointllm.py:124-129: point_backbone -> point_proj.  [B,N,C] fp32 -> [B, Np, d] bf16."""
        if self._packed_encoder is None:
            self.repack()
        B, N, _ = point_clouds.shape
        if fps_start is None:
            fps_start = self.draw_fps_start(N, B)
        fps_start = fps_start.to(self.dev)
        pts = point_clouds.to(self.dev, torch.float32)
        if enc_train is not None:          # --unfreeze_pc_encoder: train-mode encoder, activations kept for its backward
            from .encoder_train import train_forward
            tok, enc_train["saved"] = train_forward(enc_train["pk"], pts, fps_start, enc_train.get("drop_u"))
        else:
            tok = point_encoder_forward(self._packed_encoder, pts, fps_start)
        Np, T = tok.shape[1], tok.shape[2]
        out = projector_forward(self._packed_proj, tok.view(B * Np, T), save)
        return out.view(B, Np, self.cfg.hidden_size)

    def splice_positions(self, input_ids: torch.Tensor) -> torch.Tensor:
        """Validation of pointllm.py:136-164 (same ValueErrors) -> int32 [B] first point-feature position (-1: none).
        One small D2H read, like the reference's own host-side checks."""
        c = self.point_backbone_config
        n = c["point_token_len"]
        info_d = ops.point_layout(input_ids, c["point_patch_token"], c["point_start_token"], c["point_end_token"], n)
        info = info_d.cpu()
        pos = torch.full((input_ids.shape[0],), -1, dtype=torch.int32)
        self.last_splice_all = False
        for b in range(input_ids.shape[0]):
            n_start, n_end, n_patch, first_start, end_ok, consecutive, first_patch = info[b, :7].tolist()
            if n_patch == 0:
                continue                                            # pointllm.py:138-142 (dummy projection only)
            if c["mm_use_point_start_end"]:
                if n_start != n_end:
                    raise ValueError("The number of point start tokens and point end tokens should be the same.")
                if n_start > 1:
                    raise ValueError("egoscaler_b200 supports one point block per sample.")
                if not end_ok:
                    raise ValueError("The point end token should follow the point start token.")
                pos[b] = first_start + 1
            else:
                if n_patch != n:
                    raise ValueError("The number of point patch tokens should be the same as the number of point patches.")
                if not consecutive:
                    raise ValueError("The point patch tokens should be consecutive.")
                pos[b] = first_patch
        self.last_splice_all = bool((pos >= 0).all())       # host-side fact the trainer's sparse gradient exchange sizes its buffers by
        return pos.to(self.dev)

    @staticmethod
    def _kmask(attention_mask: Optional[torch.Tensor], dev) -> Optional[torch.Tensor]:
        if attention_mask is None:
            return None
        return attention_mask.to(dev).to(torch.uint8).contiguous()

    def hidden_states(self, input_ids, attention_mask=None, point_clouds=None, fps_start=None, cache: Optional[KVCache] = None,
                      saved=None, proj_saved=None, splice_pos=None, pos_dev=None, enc_train=None, out_rows=None):
        """Everything up to (not including) the final norm.  Returns (hidden [B*L, d], splice_pos)."""
        input_ids = input_ids.to(self.dev)
        B, L = input_ids.shape
        pos0 = cache.length if (cache is not None and pos_dev is None) else 0
        feats = None
        if (L != 1 or self.training) and point_clouds is not None:       # pointllm.py:112
            if splice_pos is None:                                       # host-side checks first: the (async)
                splice_pos = self.splice_positions(input_ids)            # encoder launches then queue without a sync
            feats = self.point_features(point_clouds, fps_start, proj_saved, enc_train)
        x = ops.embed_splice(input_ids, self.model.embed_tokens.weight.detach(), feats, splice_pos if feats is not None else None,
                             self.point_backbone_config["point_token_len"])
        h = self.trunk.forward(x, B, L, self._kmask(attention_mask, self.dev), pos0, cache, saved, pos_dev, out_rows=out_rows)
        if cache is not None:
            cache.length = cache.length + L
        return h, splice_pos

    def logits_from_hidden(self, h: torch.Tensor, decode: bool = False) -> torch.Tensor:
        """final RMSNorm + lm_head (pointllm.py:227-228) -> [rows, V] view of a [rows, v_pad] bf16 buffer."""
        skinny = decode and h.shape[0] <= ops.SKINNY_MAX_M       # cached single-token step: stream the tiled copy
        stat = getattr(self.trunk, "decode_stat", None) if skinny else None     # final norm folded into lm_head (llm.py)
        fresh = False
        if skinny and (self._lm_head_tiled is None or self._lm_head_folded != (stat is not None)):
            fresh = True                                    # written by the kernel just queued: no early weight prefetch this time
            self._lm_head_folded = stat is not None
            self._lm_head_tiled = ops.retile_weight(self.lm_head.weight.detach(),
                                                    self.model.norm.weight.detach() if self._lm_head_folded else None)
        buf = torch.empty(h.shape[0], self.v_pad, device=h.device, dtype=torch.bfloat16)
        out = buf[:, :self.cfg.vocab_size]
        if stat is not None:
            ops.gemm_skinny(h, self._lm_head_tiled, out=out, weights_const=not fresh, row_scale=stat, eps=self.cfg.rms_norm_eps)
            return out
        hn = ops.rmsnorm(h, self.model.norm.weight.detach(), self.cfg.rms_norm_eps)
        if skinny:
            ops.gemm_skinny(hn, self._lm_head_tiled, out=out, weights_const=True)
        else:
            ops.gemm(hn, self.lm_head.weight.detach(), out=out)
        return out

    # ------------------------------------------------------------------ forward / generate (reference signatures)
    def forward(self, input_ids=None, attention_mask=None, past_key_values: Optional[KVCache] = None, inputs_embeds=None,
                labels=None, use_cache=None, output_attentions=None, output_hidden_states=None, point_clouds=None,
                return_dict=None, fps_start=None, drop_u=None) -> CausalLMOutput:
        """model_arch.py:53-75 -> pointllm.py:195-253.  Logits for all positions, bf16 (what autocast gives the
        reference on GPU).  ``past_key_values`` follows the parent class' semantics.

        With autograd enabled (and no KV cache) the logits carry a ``grad_fn``: ``loss.backward()`` on anything computed
        from them runs the hand-written CUDA backward and fills ``.grad`` of the trainable parameters, so the reference's
        loop body (train.py:165-184) and any ``torch.optim`` / DeepSpeed optimizer work unchanged (autograd.py)."""
        if inputs_embeds is not None or labels is not None:
            raise NotImplementedError("inputs_embeds / labels are never passed by the reference entry points")
        self.sync_derived()
        cache = past_key_values
        B, L = input_ids.shape
        if torch.is_grad_enabled() and cache is None and not use_cache:
            names = tuple(self.trainable_names())
            if names:
                from .autograd import TrajVLMFunction
                named = dict(self.named_parameters())
                logits = TrajVLMFunction.apply(self, input_ids, attention_mask, point_clouds, fps_start, drop_u, names,
                                               *[named[n] for n in names])
                return CausalLMOutput(logits=logits, past_key_values=None)
        with torch.no_grad():
            if use_cache and cache is None:
                cache = KVCache.allocate(self.cfg, B, max(self.cfg.max_position_embeddings, L), self.dev)
            h, _ = self.hidden_states(input_ids, attention_mask, point_clouds, fps_start, cache)
            logits = self.logits_from_hidden(h).view(B, L, self.cfg.vocab_size)
        return CausalLMOutput(logits=logits, past_key_values=cache)

    __call__ = forward
