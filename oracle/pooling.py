"""CPU restatement of the reference embedders' pool + L2-normalise epilogue.
TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

All arithmetic is done in float32 on numpy arrays, the precision the reference reaches on its
``mean`` path (``attention_mask.float()`` promotes the product, base.py:375) and an upper bound on
what it does on the half-precision ``cls``/``last`` paths (tests compare with half-ulp tolerance).
"""
from __future__ import annotations

import numpy as np


def pool_cls(hidden: np.ndarray, mask: np.ndarray | None = None) -> np.ndarray:
    """``last_hidden_state[:, 0]`` -- Nexus/inference/embedder/text_retrieval/base.py:372-373,
    multimodal_retrieval/generic.py:105-106."""
    return hidden[:, 0].astype(np.float32)


def pool_mean(hidden: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """``sum(h * mask[...,None].float(), dim=1) / mask.sum(dim=1, keepdim=True).float()`` --
    base.py:374-377, generic.py:107-110.  A row whose mask is all zero divides 0/0 -> NaN, as in
    the reference."""
    m = mask.astype(np.float32)
    s = (hidden.astype(np.float32) * m[:, :, None]).sum(axis=1, dtype=np.float32)
    d = m.sum(axis=1, keepdims=True, dtype=np.float32)
    with np.errstate(invalid="ignore", divide="ignore"):
        return (s / d).astype(np.float32)


def pool_last_token(hidden: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """Decoder-only pooling -- Nexus/inference/embedder/text_retrieval/decoder/base.py:330-342 and
    generic.py:111-120: if every row's last mask entry is 1 (left padding) take ``h[:, -1]``,
    else ``h[b, mask[b].sum() - 1]`` (index -1 wraps to the last position for an empty mask)."""
    b = hidden.shape[0]
    left_padding = int(mask[:, -1].sum()) == b
    if left_padding:
        return hidden[:, -1].astype(np.float32)
    seq = mask.sum(axis=1).astype(np.int64) - 1
    return hidden[np.arange(b), seq].astype(np.float32)


def pool_last_token_argmax(hidden: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """MMEB variant -- Nexus/evaluation/mmeb_v2/models.py:146-151: position of the LAST 1 in the
    mask via ``flip().argmax()``.  Differs from ``sum-1`` when the mask has interior zeros."""
    b, l = mask.shape
    last_pos = np.argmax(mask[:, ::-1] != 0, axis=1) if mask.dtype == bool else np.argmax(mask[:, ::-1], axis=1)
    col = l - last_pos - 1
    return hidden[np.arange(b), col].astype(np.float32)


def l2_normalize_torch(x: np.ndarray, eps: float = 1e-12) -> np.ndarray:
    """``torch.nn.functional.normalize(x, p=2, dim=-1)`` = ``x / max(||x||_2, eps)`` -- call sites
    base.py:334-335, decoder/base.py:309-310, generic.py:124-125, mmeb_v2/models.py:137-138."""
    x = x.astype(np.float32)
    n = np.sqrt((x.astype(np.float64) ** 2).sum(axis=-1, keepdims=True)).astype(np.float32)
    return (x / np.maximum(n, np.float32(eps))).astype(np.float32)


def l2_normalize_numpy(x: np.ndarray) -> np.ndarray:
    """``x / np.linalg.norm(x, axis=-1, keepdims=True)`` -- the ONNX/TRT host path,
    base.py:672-673,711-712,750-751: no epsilon, a zero row becomes NaN."""
    x = x.astype(np.float32)
    with np.errstate(invalid="ignore", divide="ignore"):
        return x / np.linalg.norm(x, axis=-1, keepdims=True)


_POOLERS = {
    "cls": pool_cls,
    "mean": pool_mean,
    "last": pool_last_token,
    "last_token": pool_last_token,
    "last_argmax": pool_last_token_argmax,
}


def pool_and_normalize(hidden: np.ndarray, mask: np.ndarray, pooling: str = "cls", normalize: bool = True,
                       eps: float = 1e-12) -> np.ndarray:
    """pooling() followed by F.normalize, as at base.py:333-335."""
    if pooling not in _POOLERS:
        # base.py:378-379 raises NotImplementedError for anything else
        raise NotImplementedError(f"pooling method {pooling} not implemented")
    x = _POOLERS[pooling](hidden, mask)
    return l2_normalize_torch(x, eps) if normalize else x
