"""Low-rank gradient math shared by rankDAD and powerSGD (reference-precision, any device).

This is the plain-PyTorch statement of the math.  The file-mode engines use it directly,
the sm_100a kernels in ``csrc/lowrank.cu`` are tested against it, and with ``dtype=
torch.float64`` it is the oracle of SURVEY §4.2.

The reference only names the engines (``comps/__init__.py:13-16``) and their tunables
(``dad_reduction_rank``, ``dad_num_pow_iters``, ``dad_tol`` – ``compspec.json:236-238``);
their implementation lives in the absent ``coinstac_dinunet`` package, so the algorithms
are defined here from the papers' math:

* rankDAD ("distributed auto-differentiation"): a Linear layer's weight gradient is
  ``G = Δᵀ·A`` with ``Δ`` the output deltas ``[N,out]`` and ``A`` the input activations
  ``[N,in]``.  ``G`` is never formed; a block power (subspace) iteration that only touches
  ``Δ`` and ``A`` produces rank-``r`` factors ``(Δ_r [r,out], A_r [r,in])`` with
  ``G ≈ Δ_rᵀ·A_r``.
* powerSGD (Vogels et al.): ``P = M·Q``, orthogonalise ``P``, ``Q = Mᵀ·P``,
  ``M̂ = P·Qᵀ``, error feedback ``E = M − M̂``.
"""
from __future__ import annotations

import torch


def orthogonalize(P: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """Orthonormalise the columns of ``P [n,r]`` (modified Gram-Schmidt, in fp32+).

    Columns whose residual norm collapses (rank-deficient input) are zeroed rather than
    blown up, so a true rank < r gradient still reconstructs exactly.
    """
    P = P.clone()
    n, r = P.shape
    scale = P.norm() + eps
    # A linearly dependent column leaves a residual of ~machine-epsilon times its original size,
    # not zero; normalising that noise would destroy orthogonality.  "Collapsed" is therefore judged
    # relative to the column's norm before any projection (same rule as csrc/lowrank.cu).
    rel = 1e-4 if P.dtype in (torch.float32, torch.bfloat16, torch.float16) else 1e-10
    orig = P.norm(dim=0)
    for i in range(r):
        col = P[:, i]
        nrm = col.norm()
        if nrm <= max(eps * scale, rel * orig[i]):
            P[:, i] = 0
            continue
        col = col / nrm
        P[:, i] = col
        if i + 1 < r:
            rest = P[:, i + 1:]
            rest -= torch.outer(col, col @ rest)
    return P


def _orth_fast(P: torch.Tensor) -> torch.Tensor:
    """``orthogonalize`` for the hot path: one launch of the batched sm_100a kernel
    (``dnb_orthogonalize_batch``) when ``P`` is an fp32 CUDA matrix, the Python loop otherwise."""
    if P.is_cuda and P.dtype == torch.float32 and P.shape[1] <= 64:
        from ..ops import native
        if native.available():
            from ..parallel.lowrank_fused import orthogonalize_batch
            Q = P.contiguous().clone()
            orthogonalize_batch([Q])
            return Q
    if (not P.is_cuda) and P.dtype == torch.float32 and P.dim() == 2 and P.shape[1] <= 1024:
        # host fast path: the same modified Gram-Schmidt in native code (fp64 accumulation); the Python column
        # loop below was 40 % of a simulated file-mode rankDAD run
        from ..ops import native
        if native.available(cpu_only=True) and hasattr(native.lib(), "dnb_cpu_orthogonalize"):
            import ctypes
            Q = P.contiguous().clone()
            rc = native.lib().dnb_cpu_orthogonalize(ctypes.c_void_p(Q.data_ptr()), ctypes.c_longlong(Q.shape[0]),
                                                    ctypes.c_int(Q.shape[1]), ctypes.c_float(1e-8), ctypes.c_float(1e-4))
            if rc == 0:
                return Q
    return orthogonalize(P)


_SEEDED = {}


def _seeded_normal(shape, seed, device, dtype):
    """Seeded start matrix, identical on every site.  Drawn on the CPU generator (device independent);
    device copies are cached — they are constants of (shape, seed), callers never modify them in place —
    which also keeps the low-rank step free of host-to-device copies (CUDA-graph capturable)."""
    device = torch.device(device)
    key = (tuple(int(x) for x in shape), int(seed), device.type, device.index, dtype)
    hit = _SEEDED.get(key)
    if hit is not None:
        return hit
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed) % (2 ** 31 - 1))
    v = torch.randn(*shape, generator=g, dtype=torch.float32).to(device=device, dtype=dtype)
    if device.type == "cuda" and len(_SEEDED) < 4096:
        _SEEDED[key] = v
    return v


# --------------------------------------------------------------------------- rankDAD
def dad_compress(delta: torch.Tensor, act: torch.Tensor, rank: int, num_pow_iters: int = 5,
                 tol: float = 1e-3, seed: int = 0):
    """Rank-``rank`` factors of ``G = deltaᵀ·act`` without forming ``G``.

    delta ``[N,out]``, act ``[N,in]`` → ``(Δ_r [rank,out], A_r [rank,in], iters_used)`` with
    ``G ≈ Δ_rᵀ·A_r``.  Exact when ``N <= rank`` (the factors *are* the data, zero-padded)
    and when ``rank(G) <= rank`` after one iteration.

    Iteration (block power method on ``GᵀG``):  ``U = orth(Δᵀ(A V))``, ``V = Aᵀ(Δ U)``;
    stops after ``num_pow_iters`` rounds or when the captured energy ``‖V‖_F`` changes by
    less than ``tol`` relatively (``dad_tol``).
    """
    N, n_out = delta.shape
    n_in = act.shape[1]
    r = int(rank)
    dt = torch.promote_types(delta.dtype, torch.float32)
    delta, act = delta.to(dt), act.to(dt)
    if N <= r:
        d = torch.zeros(r, n_out, dtype=dt, device=delta.device)
        a = torch.zeros(r, n_in, dtype=dt, device=delta.device)
        d[:N], a[:N] = delta, act
        return d, a, 0
    V = _seeded_normal((n_in, r), seed, delta.device, dt)
    prev = None
    used = 0
    U = None
    # On a CUDA device the early exit would cost a host sync per iteration (reading ‖V‖), so the
    # device path always runs ``num_pow_iters`` rounds — a superset of the work, never less accurate.
    check_tol = not delta.is_cuda
    for it in range(max(int(num_pow_iters), 1)):
        U = _orth_fast(delta.t() @ (act @ V))           # [out, r]
        V = act.t() @ (delta @ U)                        # [in, r]  (= Gᵀ U, carries σ)
        used = it + 1
        if check_tol:
            energy = float(V.norm())
            if prev is not None and abs(energy - prev) <= tol * max(energy, 1e-30):
                break
            prev = energy
    return U.t().contiguous(), V.t().contiguous(), used


def dad_reduce(deltas, acts, rank: int, num_pow_iters: int = 5, tol: float = 1e-3, seed: int = 0,
               average: bool = True):
    """Aggregator side: concatenate every site's factors and re-compress to ``rank``.

    ``Σ_s Δ_sᵀ A_s = cat(Δ)ᵀ cat(A)`` so the same ``dad_compress`` applies to the stacked
    ``[S·r, ·]`` factors.  ``average`` divides by the number of sites (unweighted mean over
    sites, like dSGD — SURVEY §2.3 E2).
    """
    S = len(deltas)
    d = torch.cat(list(deltas), 0)
    a = torch.cat(list(acts), 0)
    dr, ar, used = dad_compress(d, a, rank, num_pow_iters, tol, seed)
    if average:
        dr = dr / float(S)
    return dr, ar, used


def dad_reconstruct(delta_r: torch.Tensor, act_r: torch.Tensor) -> torch.Tensor:
    """``G = Δ_rᵀ·A_r`` → ``[out,in]``."""
    return delta_r.t() @ act_r


# --------------------------------------------------------------------------- powerSGD
def psgd_init_q(shape_2d, rank: int, seed: int, device="cpu", dtype=torch.float32):
    """Identical ``Q [m,r]`` on every site (seeded, then orthonormalised)."""
    n, m = shape_2d
    return orthogonalize(_seeded_normal((m, int(rank)), seed, device, dtype))


def psgd_p(M: torch.Tensor, Q: torch.Tensor) -> torch.Tensor:
    return M @ Q


def psgd_q(M: torch.Tensor, P_hat: torch.Tensor) -> torch.Tensor:
    return M.t() @ P_hat


def psgd_finish(M: torch.Tensor, P_hat: torch.Tensor, Q_mean: torch.Tensor):
    """Returns ``(M̂, E)`` with ``M̂ = P̂·Q̄ᵀ`` and the local residual ``E = M − M̂``."""
    M_hat = P_hat @ Q_mean.t()
    return M_hat, M - M_hat


def psgd_oracle_step(Ms, Qs, rank):
    """All-sites-in-one-place restatement of one powerSGD step (test oracle).

    ``Ms``: list over sites of the (gradient + error) matrix.  Returns the common ``M̂``,
    the per-site residuals and the new shared ``Q``.
    """
    S = len(Ms)
    P = sum(M @ Q for M, Q in zip(Ms, Qs)) / S
    P_hat = orthogonalize(P)
    Q = sum(M.t() @ P_hat for M in Ms) / S
    M_hat = P_hat @ Q.t()
    return M_hat, [M - M_hat for M in Ms], Q
