"""FP8 KV-cache scaling factors: the JSON file of `--quantization-param-path` and the checkpoint
parameter names — host side of SURVEY §8(f) rank 1.  Same function names, arguments, return values
and failure behaviour as the reference (python/sglang/srt/model_loader/weight_utils.py:670-857); the
reference validates the document with pydantic models, this is a plain-Python check of the same rules.
Pinned to the reference's own functions run on its own fixtures (tests/golden/kv_cache_scales.pt)."""
from __future__ import annotations

import json
import logging
from typing import Iterable, Optional, Tuple

logger = logging.getLogger(__name__)
_warned = set()


def _warn_once(msg: str) -> None:
    if msg not in _warned:
        _warned.add(msg)
        logger.warning(msg)


class _BadScales(ValueError):
    pass


def _rank_maps(doc, model_type, tp_size, num_hidden_layers):
    """{tp rank: {layer: scale}} of a scales document, or _BadScales naming the rule it breaks
    (QuantParamSchema / KVCacheQuantSchema, weight_utils.py:737-808)."""
    if not isinstance(doc, dict) or not isinstance(doc.get("kv_cache"), dict):
        raise _BadScales("no kv_cache section")
    if "model_type" not in doc:
        raise _BadScales("no model_type field")
    kv = doc["kv_cache"]
    if kv.get("dtype") != "float8_e4m3fn":
        raise _BadScales(f"scaling factors are for KV cache dtype {kv.get('dtype')!r}, not float8_e4m3fn")
    if model_type is not None and doc["model_type"] != model_type:
        raise _BadScales(f"model type is {model_type} but the scaling factors belong to {doc['model_type']}")
    raw = kv.get("scaling_factor")
    if not isinstance(raw, dict):
        raise _BadScales("no scaling_factor map")
    try:
        maps = {int(r): {int(l): float(s) for l, s in layers.items()} for r, layers in raw.items()}
    except (TypeError, ValueError, AttributeError) as e:
        raise _BadScales(f"scaling_factor is not {{rank: {{layer: float}}}}: {e}") from None
    if len(maps) != tp_size:
        raise _BadScales(f"document has TP size {len(maps)}, the engine runs TP size {tp_size}")
    for r, layers in maps.items():
        if len(layers) != num_hidden_layers:
            raise _BadScales(f"rank {r}: expected {num_hidden_layers} layers, got {len(layers)}")
    for r in range(tp_size):
        if r not in maps:
            raise _BadScales(f"no scales for TP rank {r}")
    return maps


def kv_cache_scales_loader(filename: str, tp_rank: int, tp_size: int, num_hidden_layers: int,
                           model_type: Optional[str]) -> Iterable[Tuple[int, float]]:
    """(layer index, scaling factor) pairs of this TP rank.  Any problem with the file — missing,
    not JSON, wrong dtype / model type / TP size / layer set — is logged and answered with an empty
    list, which leaves every layer at scale 1.0 (weight_utils.py:811-857)."""
    try:
        with open(filename) as f:
            maps = _rank_maps(json.load(f), model_type, tp_size, num_hidden_layers)
        mine = maps[tp_rank]
        for i in range(num_hidden_layers):
            if i not in mine:
                raise _BadScales(f"no KV cache scale for layer {i} in TP rank {tp_rank}")
        return mine.items()
    except FileNotFoundError:
        logger.error("File or directory '%s' not found.", filename)
    except json.JSONDecodeError:
        logger.error("Error decoding JSON in file '%s'.", filename)
    except Exception as e:  # noqa: BLE001 - the reference swallows every other error too
        logger.error("An error occurred while reading '%s': %s", filename, e)
    logger.warning("Defaulting to KV cache scaling factors = 1.0 for all layers in TP rank %d "
                   "as an error occurred during loading.", tp_rank)
    return []


def maybe_remap_kv_scale_name(name: str, params_dict: dict) -> Optional[str]:
    """Checkpoint name of a KV-cache scale -> the attention layer's parameter name
    (weight_utils.py:670-733): `.kv_scale` (deprecated single scale) -> `.attn.k_scale`;
    `.k_scale` / `.v_scale` -> `.attn.k_scale` / `.attn.v_scale`, also from ModelOpt's
    `.self_attn.k_proj.k_scale` / `.self_attn.v_proj.v_scale`.  None when the model has no such
    parameter (the scale is skipped); any other name is returned unchanged."""
    if name.endswith(".kv_scale"):
        _warn_once("DEPRECATED. Found kv_scale in the checkpoint: remapped to k_scale and "
                   "duplicated to v_scale; use separate k_scale and v_scale tensors")
        target = name.replace(".kv_scale", ".attn.k_scale")
    else:
        for kind in ("k", "v"):
            suffix = f".{kind}_scale"
            if name.endswith(suffix):
                modelopt = f".self_attn.{kind}_proj{suffix}"
                if modelopt in name:
                    target = name.replace(modelopt, f".self_attn.attn{suffix}")
                else:
                    target = name.replace(suffix, f".attn{suffix}")
                break
        else:
            return name
    if target not in params_dict:
        _warn_once(f"Found a KV cache scale in the checkpoint (e.g. {name}) but the model has no "
                   f"parameter {target}; it is not loaded")
        return None
    return target


def load_kv_cache_scales(attn_layers, quantization_param_path: str, tp_rank: int, tp_size: int,
                         model_type: Optional[str]) -> int:
    """What the reference models' load_kv_cache_scales does (models/llama.py:359-378): the file's
    per-layer factor becomes BOTH k_scale and v_scale of layer i's RadixAttention.  Also fills the
    host copies k_scale_float / v_scale_float the B200 backend reads (no device sync per step).
    `attn_layers[i]` is layer i's attention layer.  Returns the number of layers set (0 = file
    unusable, scales stay at their defaults)."""
    import torch

    n = 0
    for layer_idx, factor in kv_cache_scales_loader(quantization_param_path, tp_rank, tp_size,
                                                    len(attn_layers), model_type):
        attn = attn_layers[layer_idx]
        if not hasattr(attn, "k_scale"):
            raise RuntimeError("Self attention has no KV cache scaling factor attribute!")
        for which in ("k_scale", "v_scale"):
            cur = getattr(attn, which)
            if isinstance(cur, torch.nn.Parameter):     # created by a KV-cache quant method
                cur.data.fill_(factor)
            else:
                setattr(attn, which, factor)
            setattr(attn, which + "_float", float(factor))
        n += 1
    return n
