"""TEST INFRASTRUCTURE ONLY - torch restatement of the four quaternion helpers the Go2 envs import
(reference: genesis/utils/geom.py:430-440 inv_quat, :718-775 quat_to_xyz, :990-1021 transform_quat_by_quat,
:1052-1086 transform_by_quat)."""
import torch

EPS = 1.1920929e-07


def inv_quat(quat):
    q = quat.clone()
    q[..., 1:] = -q[..., 1:]
    return q


def transform_quat_by_quat(v, u):
    """quatmul(u, v), normalised."""
    w1, x1, y1, z1 = u.unbind(-1)
    w2, x2, y2, z2 = v.unbind(-1)
    out = torch.stack(
        [
            w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2,
            w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
            w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
            w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2,
        ],
        dim=-1,
    )
    return out / torch.linalg.vector_norm(out, dim=-1, keepdim=True)


def transform_by_quat(v, quat):
    qw, qx, qy, qz = quat[..., :1], quat[..., 1:2], quat[..., 2:3], quat[..., 3:]
    ww, wx, wy, wz = qw * qw, qw * qx, qw * qy, qw * qz
    xx, xy, xz = qx * qx, qx * qy, qx * qz
    yy, yz, zz = qy * qy, qy * qz, qz * qz
    vs = v / (ww + xx + yy + zz)
    vx, vy, vz = vs[..., :1], vs[..., 1:2], vs[..., 2:]
    return torch.cat(
        [
            vx * (xx + ww - yy - zz) + vy * (2.0 * xy - 2.0 * wz) + vz * (2.0 * xz + 2.0 * wy),
            vx * (2.0 * wz + 2.0 * xy) + vy * (ww - xx + yy - zz) + vz * (2.0 * yz - 2.0 * wx),
            vx * (2.0 * xz - 2.0 * wy) + vy * (2.0 * wx + 2.0 * yz) + vz * (ww - xx - yy + zz),
        ],
        dim=-1,
    )


def quat_to_xyz(quat, rpy=False, degrees=False):
    qw, qx, qy, qz = quat[..., 0], quat[..., 1], quat[..., 2], quat[..., 3]
    ww, wx, wy, wz = qw * qw, qw * qx, qw * qy, qw * qz
    xx, xy, xz = qx * qx, qx * qy, qx * qz
    yy, yz, zz = qy * qy, qy * qz, qz * qz
    if rpy:
        sinp, sinrcosp, sinycosp = wy - xz, wx + yz, wz + xy
    else:
        sinp, sinrcosp, sinycosp = xz + wy, wx - yz, wz - xy
    cosrcosp = (ww - xx - yy + zz) / 2
    cosycosp = (ww + xx - yy - zz) / 2
    cosp = torch.sqrt(cosycosp**2 + sinycosp**2)
    x = torch.atan2(sinrcosp, cosrcosp)
    y = torch.atan2(sinp, cosp)
    z = torch.atan2(sinycosp, cosycosp)
    mask = cosp < EPS
    alt = (wz - xy) if rpy else (wz + xy)
    z = torch.where(mask, torch.atan2(alt, (ww - xx + yy - zz) / 2), z)
    x = torch.where(mask, torch.zeros_like(x), x)
    out = torch.stack([x, y, z], dim=-1)
    return torch.rad2deg(out) if degrees else out
