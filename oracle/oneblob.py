"""OneBlob encoding -- CPU restatement (torch, differentiable).  PARITY UNPINNED.

The reference builds ``tcnn.Encoding(n_input_dims=3, {"otype": "OneBlob", "n_bins": 16}, dtype=float)``
(``model/encodings.py:61-71``) and calls it on ``inputs_flat`` (``model/scene_rep.py:259,294``).  The
arithmetic lives in tiny-cuda-nn (``include/tiny-cuda-nn/encodings/oneblob.h``), an un-vendored
dependency pinned only as git HEAD in ``requirements.txt:120`` (fallback commit
``91ee479d275d322a65726435040fc20b56b9c991``, ``README.md:99``).  It cannot be installed here and the
reference has no test vectors for it, so this file restates the published algorithm:

* quartic kernel of radius ``1/n_bins``; its CDF is
  ``cdf(t) = clamp(15/16 * u * (1 - 2/3 u^2 + 1/5 u^4) + 1/2, 0, 1)`` with ``u = t * n_bins``;
* the value of bin ``b`` for input ``x`` is ``C(edge_{b+1} - x) - C(edge_b - x)`` where
  ``C(t) = cdf(t) + cdf(t-1) + cdf(t+1)`` adds the two wrap-around images and ``edge_b = b / n_bins``;
* for a power-of-two bin count tiny-cuda-nn evaluates only left edges and takes the right edge of bin
  ``b`` from the left edge of bin ``b+1`` (the last bin takes ``C(edge_0 - x) + 1``).  That is identical
  to the direct formula for ``x`` in ``[-15/16, 31/16]``; the restatement follows the power-of-two
  variant because ``n_bins = 16`` is the only value the reference's configs use.

Output layout: ``[M, n_dims * n_bins]`` with the bins of dimension 0 first (tcnn: encoded index
``= dim * n_bins + bin``).
"""
import torch


def _cdf(t, n_bins):
    u = t * n_bins
    u2 = u * u
    u4 = u2 * u2
    return torch.clamp((15.0 / 16.0) * u * (1 - (2.0 / 3.0) * u2 + (1.0 / 5.0) * u4) + 0.5, 0.0, 1.0)


def _wrapped_cdf(t, n_bins):
    return _cdf(t, n_bins) + _cdf(t - 1.0, n_bins) + _cdf(t + 1.0, n_bins)


def oneblob(x, n_bins=16):
    """x: [M, D] float32 -> [M, D * n_bins] float32."""
    x = x.float()
    M, D = x.shape
    pow2 = (n_bins & (n_bins - 1)) == 0
    left_edges = torch.arange(n_bins, dtype=torch.float32, device=x.device) / n_bins            # [B]
    left = _wrapped_cdf(left_edges[None, None, :] - x[:, :, None], n_bins)       # [M, D, B]
    if pow2:
        right = torch.cat([left[..., 1:], left[..., :1] + 1.0], dim=-1)
    else:
        right_edges = (torch.arange(n_bins, dtype=torch.float32, device=x.device) + 1) / n_bins
        right = _wrapped_cdf(right_edges[None, None, :] - x[:, :, None], n_bins)
    return (right - left).reshape(M, D * n_bins)
