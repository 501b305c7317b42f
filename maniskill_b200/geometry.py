"""Quaternion helpers (wxyz) for the non-hot bookkeeping around the kernels.

Semantics follow mani_skill/utils/geometry/rotation_conversions.py (the functions
the controllers use: quaternion_multiply :407-421, quaternion_apply :441-461,
quaternion_to_matrix :44-73); the hot-path versions live in csrc/msik_math.cuh.
"""
from __future__ import annotations

import torch


def quat_raw_mul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    aw, ax, ay, az = a.unbind(-1)
    bw, bx, by, bz = b.unbind(-1)
    return torch.stack(
        (
            aw * bw - ax * bx - ay * by - az * bz,
            aw * bx + ax * bw + ay * bz - az * by,
            aw * by - ax * bz + ay * bw + az * bx,
            aw * bz + ax * by - ay * bx + az * bw,
        ),
        -1,
    )


def quat_standardize(q: torch.Tensor) -> torch.Tensor:
    return torch.where(q[..., :1] < 0, -q, q)


def quat_mul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    return quat_standardize(quat_raw_mul(a, b))


def quat_conj(q: torch.Tensor) -> torch.Tensor:
    return q * q.new_tensor([1.0, -1.0, -1.0, -1.0])


def quat_apply(q: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    vq = torch.cat((v.new_zeros(v.shape[:-1] + (1,)), v), -1)
    return quat_raw_mul(quat_raw_mul(q, vq), quat_conj(q))[..., 1:]


def quat_to_matrix(q: torch.Tensor) -> torch.Tensor:
    r, i, j, k = q.unbind(-1)
    s = 2.0 / (q * q).sum(-1)
    m = torch.stack(
        (
            1 - s * (j * j + k * k), s * (i * j - k * r), s * (i * k + j * r),
            s * (i * j + k * r), 1 - s * (i * i + k * k), s * (j * k - i * r),
            s * (i * k - j * r), s * (j * k + i * r), 1 - s * (i * i + j * j),
        ),
        -1,
    )
    return m.reshape(q.shape[:-1] + (3, 3))
