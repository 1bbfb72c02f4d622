"""Checkpoint -> PQ linears.

Mirrors the part of the reference's model loading that touches the PQ path:

* `safetensors_weights_iterator` (vllm/model_executor/model_loader/weight_utils.py) - stream
  `(name, tensor)` pairs out of the checkpoint's `*.safetensors` shards;
* `LlamaModel.load_weights` (vllm/model_executor/models/llama.py:369-433) - the
  `stacked_params_mapping` walk that routes `q_proj / k_proj / v_proj` into `qkv_proj` and
  `gate_proj / up_proj` into `gate_up_proj` through each parameter's `weight_loader` with its shard
  id, and everything else through the parameter's own loader;
* `AQLMConfig.from_config` on the `quantization_config` block of the HF `config.json`
  (aqlm.py:198-204) - here `PQConfig.from_config`.

An AQLM checkpoint stores, per linear, `<prefix>.codes [N, G, M]`, `<prefix>.codebooks
[M, 2^b, 1, d]` and `<prefix>.scales [N, 1, 1, 1]` - exactly the tensors `create_weights` registers,
so the weight loaders only narrow (tensor parallelism) and place (fused layers): the PARAMETERS
keep the checkpoint's layout.  `load_checkpoint` then runs `process_weights_after_loading`, which
may add private device copies next to them (codes rows padded to a 16-byte pitch for TMA, 4-bit codes
packed two to a byte against a product codebook - DESIGN.md 4); those are derived data, never loaded.

Beyond the four Llama projections the walk also places
* the vocabulary layers when the model holds PQ ones (`embed_tokens`, `lm_head` of
  vllm_pq_b200/embedding.py - a checkpoint whose config says `lm_head_quantized`): the rank's rows of
  `model.embed_tokens.{codes,codebooks,scales}` / `lm_head.*` (vocab_parallel_embedding.py:361-393);
* the experts of a quantised Mixtral block (`layers.N.block_sparse_moe.experts.E.w1|w2|w3.*` and the
  dense router `...gate.weight`), skipping experts that live on another rank
  (mixtral_quant.py:436-439).
Everything else (dense embeddings, norms, rotary caches) is outside this package's scope and is
returned as `skipped`, not loaded.
"""
from __future__ import annotations

import glob
import json
import os
import re
from typing import Dict, Iterable, Iterator, List, Set, Tuple

import torch

from .pq import PQConfig

# (fused parameter owner, checkpoint module, shard id) - llama.py:371-378
STACKED_PARAMS_MAPPING = [
    (".qkv_proj", ".q_proj", "q"),
    (".qkv_proj", ".k_proj", "k"),
    (".qkv_proj", ".v_proj", "v"),
    (".gate_up_proj", ".gate_proj", 0),
    (".gate_up_proj", ".up_proj", 1),
]

_LAYER_RE = re.compile(r"^(?:model\.)?layers\.(\d+)\.(?:self_attn|mlp)(\..+)$")
_VOCAB_RE = re.compile(r"^(?:model\.)?(embed_tokens|lm_head)\.(codes|codebooks|scales|bias)$")
_MOE_RE = re.compile(r"^(?:model\.)?layers\.(\d+)\.block_sparse_moe\.(experts\.(\d+)\.w[123]\.\w+|gate\.weight)$")
_PQ_KEYS = ("codes", "codebooks", "scales")


def quant_config_from_hf(config: Dict) -> PQConfig:
    """`config.json` (already parsed) -> PQConfig; raises ValueError if the checkpoint is not
    quantized with a method this plugin serves (aqlm / pq)."""
    qc = config.get("quantization_config")
    if not qc:
        raise ValueError("checkpoint config has no quantization_config")
    method = str(qc.get("quant_method", "")).lower()
    if method not in ("aqlm", "pq"):
        raise ValueError(f"quantization method {method!r} is not served by the PQ plugin")
    return PQConfig.from_config(qc)


def safetensors_weights_iterator(path_or_files) -> Iterator[Tuple[str, torch.Tensor]]:
    """Yield (name, tensor) from a checkpoint directory or a list of .safetensors files, shard by
    shard (tensors are materialised one at a time: a 70B AQLM checkpoint is 9-17 GB of codes)."""
    from safetensors import safe_open

    if isinstance(path_or_files, (str, os.PathLike)):
        files = sorted(glob.glob(os.path.join(os.fspath(path_or_files), "*.safetensors")))
        if not files:
            raise FileNotFoundError(f"no *.safetensors under {path_or_files}")
    else:
        files = list(path_or_files)
    for f in files:
        with safe_open(f, framework="pt", device="cpu") as fh:
            for name in fh.keys():
                yield name, fh.get_tensor(name)


def _target_name(name: str):
    """Checkpoint tensor name -> (layer index, '.qkv_proj.codes'-style suffix) or None."""
    m = _LAYER_RE.match(name)
    if not m:
        return None
    return int(m.group(1)), m.group(2)


def _load_vocab(name: str, tensor: torch.Tensor, stack: torch.nn.Module, params_dict, loaded: Set[str]) -> bool:
    """`model.embed_tokens.codes` / `lm_head.scales` ... -> the PQ vocabulary layer of that name, if the
    model has one.  A PQ tensor for a vocabulary layer the model keeps dense (or lacks) is an error
    only when the OTHER vocabulary layer is a PQ one - then the checkpoint and the model disagree."""
    m = _VOCAB_RE.match(name)
    if not m:
        return False
    pname = f"{m.group(1)}.{m.group(2)}"
    if pname not in params_dict:
        if m.group(2) == "bias":
            return False
        if getattr(getattr(stack, m.group(1), None), m.group(2), None) is not None:
            return True                                 # tied to the other layer's parameter: already placed
        if any(p.split(".")[0] in ("embed_tokens", "lm_head") for p in params_dict):
            raise KeyError(f"{name}: the model has no parameter {pname}")
        return False
    param = params_dict[pname]
    param.weight_loader(param, tensor)
    loaded.add(pname)
    return True


def _load_moe(name: str, tensor: torch.Tensor, stack: torch.nn.Module, params_dict, loaded: Set[str]) -> bool:
    """Mixtral expert / router tensors (mixtral_quant.py:432-446).  Returns True when the name was
    consumed: loaded, or an expert of another rank (skipped on purpose, not reported as `skipped`)."""
    m = _MOE_RE.match(name)
    if not m:
        return False
    layer, tail = int(m.group(1)), m.group(2)
    base = f"layers.{layer}.block_sparse_moe"
    if not any(p.startswith(base + ".") for p in params_dict):
        if any(".block_sparse_moe." in p for p in params_dict):
            raise KeyError(f"{name}: the model has no MoE block in layer {layer}")
        return False
    pname = f"{base}.gate_weight" if tail == "gate.weight" else f"{base}.{tail}"
    if pname not in params_dict:
        if tail == "gate.weight" or pname.endswith(".bias"):
            return False
        moe = stack.get_submodule(base)
        expert = int(m.group(3))
        if expert >= len(moe.experts) or tail.rsplit(".", 1)[-1] not in _PQ_KEYS:
            raise KeyError(f"{name}: the model has no parameter {pname}")
        return True                                     # another rank's expert
    param = params_dict[pname]
    loader = getattr(param, "weight_loader", None)
    if loader is None:                                  # the dense router: a plain parameter
        if param.shape != tensor.shape:
            raise ValueError(f"{name}: shape {tuple(tensor.shape)} != {tuple(param.shape)}")
        param.data.copy_(tensor)
    else:
        loader(param, tensor)
    loaded.add(pname)
    return True


def load_pq_linears(stack: torch.nn.Module, weights: Iterable[Tuple[str, torch.Tensor]]):
    """Route checkpoint tensors into the PQ linears of `stack` (a PQLlamaLinearStack, or any module
    with `.layers[i].{qkv_proj,o_proj,gate_up_proj,down_proj}`; optionally `.embed_tokens`, `.lm_head`
    and `.layers[i].block_sparse_moe`).

    Returns (loaded, skipped): the parameter names that received data, and the checkpoint names
    that do not belong to a PQ linear.  Raises KeyError for a tensor that names a PQ linear the
    stack does not have (layer index out of range, unknown projection) - a silent skip there would
    leave random weights in the model."""
    params_dict = dict(stack.named_parameters())
    loaded: Set[str] = set()
    skipped: List[str] = []
    for name, tensor in weights:
        tgt = _target_name(name)
        if tgt is None or "rotary_emb" in name:
            if not (_load_vocab(name, tensor, stack, params_dict, loaded)
                    or _load_moe(name, tensor, stack, params_dict, loaded)):
                skipped.append(name)
            continue
        layer, suffix = tgt
        for param_name, weight_name, shard_id in STACKED_PARAMS_MAPPING:
            if weight_name not in suffix:
                continue
            pname = f"layers.{layer}{suffix.replace(weight_name, param_name)}"
            if pname.endswith(".bias") and pname not in params_dict:
                break                                   # llama.py:409-411: extra bias, skip
            param = params_dict[pname]
            param.weight_loader(param, tensor, shard_id)
            loaded.add(pname)
            break
        else:
            pname = f"layers.{layer}{suffix}"
            if pname.endswith(".bias") and pname not in params_dict:
                continue
            if pname not in params_dict:
                if suffix.split(".")[1] in ("o_proj", "down_proj"):
                    raise KeyError(f"{name}: the stack has no parameter {pname}")
                skipped.append(name)                    # e.g. q_norm / rotary caches inside self_attn
                continue
            param = params_dict[pname]
            param.weight_loader(param, tensor)
            loaded.add(pname)
    return loaded, skipped


def load_checkpoint(stack: torch.nn.Module, checkpoint_dir: str, strict: bool = True):
    """Load every PQ linear of `stack` from an HF-layout checkpoint directory and run
    `process_weights_after_loading`.  With `strict`, every codes / codebooks / scales parameter of
    the stack must have been covered by the checkpoint."""
    loaded, skipped = load_pq_linears(stack, safetensors_weights_iterator(checkpoint_dir))
    if strict:
        want = {n for n, _ in stack.named_parameters() if n.rsplit(".", 1)[-1] in _PQ_KEYS}
        missing = sorted(want - loaded)
        if missing:
            raise KeyError(f"checkpoint does not cover {len(missing)} PQ parameters, e.g. {missing[:4]}")
    for mod in stack.modules():
        qm = getattr(mod, "quant_method", None)
        if qm is not None and hasattr(qm, "process_weights_after_loading"):
            qm.process_weights_after_loading(mod)
    return loaded, skipped


def read_hf_config(checkpoint_dir: str) -> Dict:
    with open(os.path.join(checkpoint_dir, "config.json")) as fh:
        return json.load(fh)
