"""torch versions of the four quaternion helpers the eval scripts import from genesis.utils.geom
(w, x, y, z convention; transform_quat_by_quat(v, u) == normalise(u * v))."""
import torch


def inv_quat(q):
    return torch.cat([q[..., :1], -q[..., 1:]], dim=-1)


def transform_quat_by_quat(v, u):
    uw, ux, uy, uz = u.unbind(-1)
    vw, vx, vy, vz = v.unbind(-1)
    out = torch.stack([uw * vw - ux * vx - uy * vy - uz * vz, uw * vx + ux * vw + uy * vz - uz * vy,
                       uw * vy - ux * vz + uy * vw + uz * vx, uw * vz + ux * vy - uy * vx + uz * vw], dim=-1)
    return out / out.norm(dim=-1, keepdim=True)


def transform_by_quat(v, q):
    w, xyz = q[..., :1], q[..., 1:]
    n2 = (q * q).sum(-1, keepdim=True)
    t = 2.0 * torch.cross(xyz, v, dim=-1)
    return (v * n2 + w * t * 1.0 + torch.cross(xyz, t, dim=-1) - v * (n2 - (w * w + (xyz * xyz).sum(-1, keepdim=True)))) / n2


def quat_to_xyz(q, rpy=False, degrees=False):
    w, x, y, z = q.unbind(-1)
    if rpy:
        sinp, srcp, sycp = w * y - x * z, w * x + y * z, w * z + x * y
    else:
        sinp, srcp, sycp = x * z + w * y, w * x - y * z, w * z - x * y
    crcp, cycp = (w * w - x * x - y * y + z * z) / 2, (w * w + x * x - y * y - z * z) / 2
    cosp = torch.sqrt(cycp**2 + sycp**2)
    out = torch.stack([torch.atan2(srcp, crcp), torch.atan2(sinp, cosp), torch.atan2(sycp, cycp)], dim=-1)
    return torch.rad2deg(out) if degrees else out
