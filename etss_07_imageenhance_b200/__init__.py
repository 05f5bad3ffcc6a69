"""B200-native (sm_100a) drop-in for the inference hot path of `mon`'s Zero-DCE family.

Public surface (mirrors the reference for this path only):
  MODELS                         registry with mon's ModelFactory semantics (registry.py)
  ZeroDCE_RE, ZeroDCEpp_RE       `mon` model classes (models.py) — forward(datapoint) -> dict
  ZeroDCEV                       ZeroDCE-V, core of the forward only (forward_core; SURVEY 8f-3, partial)
  SCI_RE                         upstream SCI Finetunemodel
  RUAS_RE                        upstream RUAS Network (3 IEM iterations + denoise net)
  torch.ops.b200dce.*            custom ops over the C ABI (ops.py, include/b200dce.h)
  FramePipeline, shard_range     host-fed frame scheduler / multi-GPU frame sharding (sched.py)
  DceppBatchStream               device-resident batch stream: stage 1 of batch i+1 queued behind stage 2 of batch i
  GraphedForward                 model(datapoint) of one input shape captured into a CUDA graph and replayed

There is no CPU fallback: the kernels live in lib/libb200dce.so (build: `python -m
etss_07_imageenhance_b200.build`); a missing library or a CPU tensor raises.
"""
from . import ops  # registers torch.ops.b200dce.*
from .models import (DSConv2d, DepthwiseSeparableConv2d, SCI_RE, RUAS_RE, ZeroDCE_RE, ZeroDCEV, ZeroDCEpp_RE, ZeroDCELoss,
                     remap_upstream_dcepp_keys)
from .registry import MODELS, ModelFactory, register_into_mon
from .sched import DceppBatchStream, FramePipeline, GraphedForward, gather_max_ms, gather_objects, shard_range

# Drop-in step: when `mon` imports in this interpreter, the B200 classes replace the reference's entries in
# mon.MODELS (register(..., replace=True), src/mon/core/factory.py:300-328); without `mon` this is a no-op.
MON_REGISTERED = register_into_mon()

__version__ = "0.2.0"
__all__ = ["ops", "MODELS", "ModelFactory", "register_into_mon", "ZeroDCE_RE", "ZeroDCEV", "ZeroDCEpp_RE", "SCI_RE", "RUAS_RE", "DSConv2d",
           "DepthwiseSeparableConv2d", "ZeroDCELoss", "remap_upstream_dcepp_keys", "FramePipeline", "DceppBatchStream", "GraphedForward", "shard_range",
           "gather_max_ms", "gather_objects"]
