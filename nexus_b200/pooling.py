"""Fused pool + L2-normalise epilogue over encoder last-hidden-states (kernel K0, csrc/pool.cu).

Mirrors, on CUDA tensors:
  * ``BaseEmbedder.pooling`` + ``F.normalize``      Nexus/inference/embedder/text_retrieval/base.py:333-335,355-379
  * ``BaseDecoderEmbedder.pooling``                 .../text_retrieval/decoder/base.py:309-310,330-342
  * ``MultimodalEmbedder._pool_hidden_states``      .../multimodal_retrieval/generic.py:104-126
  * ``MMEBEmbeddingModel._pooling_last``            Nexus/evaluation/mmeb_v2/models.py:136-151
"""
from __future__ import annotations

from typing import Optional

from . import _lib as L
from .flat import _stream_ptr

_POOLING = {"cls": L.POOL_CLS, "mean": L.POOL_MEAN, "last": L.POOL_LAST, "last_token": L.POOL_LAST,
            "last_argmax": L.POOL_LAST_ARGMAX}


def pool_and_normalize(last_hidden_state, attention_mask=None, pooling_method: str = "cls", normalize: bool = True,
                       eps: float = 1e-12, out_dtype=None, out=None):
    """``F.normalize(pooling(last_hidden_state, attention_mask), dim=-1)`` in one kernel.

    ``last_hidden_state`` [B, L, H] CUDA tensor (fp32 / fp16 / bf16), ``attention_mask`` [B, L] (any integer
    or bool dtype).  Output [B, H]: fp32 for ``mean`` (the reference promotes through ``mask.float()``,
    base.py:375) and the input dtype otherwise, unless ``out_dtype`` says differently (e.g. torch.bfloat16
    to write straight into an index shard).  ``eps <= 0`` selects the reference's host-numpy form without
    the clamp (base.py:672-673).  Unknown pooling methods raise NotImplementedError like base.py:378-379."""
    import torch
    if pooling_method not in _POOLING:
        raise NotImplementedError(f"pooling method {pooling_method} not implemented")
    h = last_hidden_state
    if not h.is_cuda:
        raise RuntimeError("pool_and_normalize runs on CUDA tensors only (no CPU fallback)")
    if h.dim() != 3:
        raise ValueError(f"last_hidden_state must be [B, L, H], got {tuple(h.shape)}")
    codes = {torch.float32: L.F32, torch.float16: L.F16, torch.bfloat16: L.BF16}
    if h.dtype not in codes:
        h = h.float()
    h = h.contiguous()
    B, Lq, H = h.shape
    mask_ptr = 0
    m = None
    if attention_mask is not None:
        m = attention_mask.to(device=h.device, dtype=torch.int64).contiguous()
        if tuple(m.shape) != (B, Lq):
            raise ValueError(f"attention_mask must be [B, L]={B, Lq}, got {tuple(m.shape)}")
        mask_ptr = m.data_ptr()
    elif pooling_method != "cls":
        raise ValueError(f"pooling method {pooling_method} needs attention_mask")
    if out_dtype is None:
        out_dtype = torch.float32 if pooling_method == "mean" else h.dtype
    if out is None:
        out = torch.empty((B, H), dtype=out_dtype, device=h.device)
    elif tuple(out.shape) != (B, H) or not out.is_cuda or not out.is_contiguous():
        raise ValueError("out must be a contiguous CUDA tensor of shape [B, H]")
    dev = h.device.index
    L.check(L.load().nxf_pool_normalize(h.data_ptr(), codes[h.dtype], mask_ptr, B, Lq, H, _POOLING[pooling_method],
                                        1 if normalize else 0, float(eps), out.data_ptr(), codes[out.dtype], dev,
                                        _stream_ptr(dev)))
    return out


def pooling(last_hidden_state, attention_mask=None, pooling_method: str = "cls"):
    """The reference's ``pooling()`` alone (no normalisation)."""
    return pool_and_normalize(last_hidden_state, attention_mask, pooling_method, normalize=False)
