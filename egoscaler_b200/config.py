"""Model dimensions for the trajectory VLM (reference: HF ``config.json`` of the PointLLM
checkpoint read at pointllm/model/pointllm.py:32-59 plus the PointBERT yaml files
pointllm/model/pointbert/PointTransformer_8192point_2layer.yaml:1-16 and
PointTransformer_base_8192point.yaml:1-13)."""
from __future__ import annotations

from dataclasses import dataclass, field, asdict
from typing import Tuple

# Vocabulary layout after builder.py:33-46 (add_trajectory_token) on a 32000-token LLaMA vocab.
POINT_PATCH_TOKEN_ID = 32000
POINT_START_TOKEN_ID = 32001
POINT_END_TOKEN_ID = 32002
TS_TOKEN_ID = 32003        # <ts>
TSEP_TOKEN_ID = 32004      # <tsep>
TE_TOKEN_ID = 32005        # <te>
TRAJ_TOKEN_BASE = 32006    # <p0>
VOCAB_SIZE = 32262


@dataclass
class PointEncoderConfig:
    name: str = "PointTransformer_8192point_2layer"
    trans_dim: int = 384
    depth: int = 12
    num_heads: int = 6
    group_size: int = 32
    num_group: int = 512
    encoder_dims: int = 256
    point_dims: int = 6
    projection_hidden_dim: Tuple[int, ...] = (1024, 2048)
    use_max_pool: bool = False
    npoints: int = 8192

    @property
    def point_token_len(self) -> int:
        return 1 if self.use_max_pool else self.num_group + 1

    @staticmethod
    def v12() -> "PointEncoderConfig":
        return PointEncoderConfig()

    @staticmethod
    def base() -> "PointEncoderConfig":
        return PointEncoderConfig(name="PointTransformer_base_8192point", trans_dim=1152, num_heads=12,
                                  group_size=48, encoder_dims=512, projection_hidden_dim=())


@dataclass
class TrajVLMConfig:
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    vocab_size: int = VOCAB_SIZE
    rms_norm_eps: float = 1e-6
    rope_theta: float = 10000.0
    max_position_embeddings: int = 2048
    pad_token_id: int = 0
    bos_token_id: int = 1
    eos_token_id: int = 2
    mm_use_point_start_end: bool = True
    point: PointEncoderConfig = field(default_factory=PointEncoderConfig)

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @staticmethod
    def tiny() -> "TrajVLMConfig":
        """BASELINE.json configs[0]: 2 layers, d=256, 4 heads x 64, MLP 688."""
        return TrajVLMConfig(hidden_size=256, intermediate_size=688, num_hidden_layers=2, num_attention_heads=4)

    @staticmethod
    def tiny_hd128() -> "TrajVLMConfig":
        """Tiny trunk with the full-size head width (2 heads x 128): the parity-test size that runs the same head_dim-128
        tcgen05 attention / decode-attention kernels as the full-size config."""
        return TrajVLMConfig(hidden_size=256, intermediate_size=688, num_hidden_layers=2, num_attention_heads=2)

    @staticmethod
    def tiny_base() -> "TrajVLMConfig":
        """Tiny LLM + the base PointBERT encoder (SURVEY 8d config E at parity-test size)."""
        return TrajVLMConfig(hidden_size=256, intermediate_size=688, num_hidden_layers=2, num_attention_heads=4,
                             point=PointEncoderConfig.base())

    @staticmethod
    def full() -> "TrajVLMConfig":
        """LLaMA-7B / Vicuna dims used by RunsenXu/PointLLM_7B_v1.2 (train.py:321)."""
        return TrajVLMConfig()

    @staticmethod
    def full_base() -> "TrajVLMConfig":
        return TrajVLMConfig(point=PointEncoderConfig.base())

    def to_dict(self):
        return asdict(self)


def config_from_hf(hf) -> TrajVLMConfig:
    """The HF / PointLLM config object the reference passes around (``AutoConfig.from_pretrained(model_name)``,
    builder.py:14; fields read at pointllm.py:32-59) -> TrajVLMConfig.  ``vocab_size`` is taken as given: build the
    config with the post-``add_trajectory_token`` size (32262) - see ``TrajPointLLMForCausalLM.resize_token_embeddings``."""
    if hf is None:
        return TrajVLMConfig.full()
    g = lambda k, dflt: getattr(hf, k, dflt) if not isinstance(hf, dict) else hf.get(k, dflt)
    name = g("point_backbone_config_name", "PointTransformer_8192point_2layer")
    if "base" in str(name):
        point = PointEncoderConfig.base()
    else:
        point = PointEncoderConfig.v12()
    heads = int(g("num_attention_heads", 32))
    kv = g("num_key_value_heads", heads)
    if kv is not None and int(kv) != heads:
        raise ValueError("egoscaler_b200 implements multi-head attention (LLaMA-7B / Vicuna): num_key_value_heads must equal "
                         "num_attention_heads")
    rope = g("rope_theta", None)
    if rope is None:
        rp = g("rope_parameters", None) or {}
        rope = rp.get("rope_theta", 10000.0) if isinstance(rp, dict) else 10000.0
    pad = g("pad_token_id", 0)
    return TrajVLMConfig(hidden_size=int(g("hidden_size", 4096)), intermediate_size=int(g("intermediate_size", 11008)),
                         num_hidden_layers=int(g("num_hidden_layers", 32)), num_attention_heads=heads,
                         vocab_size=int(g("vocab_size", VOCAB_SIZE)), rms_norm_eps=float(g("rms_norm_eps", 1e-6)),
                         rope_theta=float(rope), max_position_embeddings=int(g("max_position_embeddings", 2048)),
                         pad_token_id=int(0 if pad is None else pad), bos_token_id=int(g("bos_token_id", 1) or 1),
                         eos_token_id=int(g("eos_token_id", 2) or 2),
                         mm_use_point_start_end=bool(g("mm_use_point_start_end", True)), point=point)
