"""Euler / quaternion helpers.  Same functions as the reference's `utils/transform.py`, written
on plain numpy (Pinocchio is not available): ZYX Euler angles stored as (yaw, pitch, roll),
quaternions as (x, y, z, w) like `pin.Quaternion.coeffs()`."""
import numpy as np


def rpy_to_matrix(rpy: np.ndarray) -> np.ndarray:
    """pin.rpy.rpyToMatrix: R = Rz(yaw) Ry(pitch) Rx(roll), argument order (roll, pitch, yaw)."""
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    return np.array([
        [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
        [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
        [-sp, cp * sr, cp * cr],
    ])


def matrix_to_rpy(R: np.ndarray) -> np.ndarray:
    """pin.rpy.matrixToRpy: (roll, pitch, yaw) with pitch in [-pi/2, pi/2]."""
    m = np.hypot(R[2, 1], R[2, 2])
    p = np.arctan2(-R[2, 0], m)
    if abs(abs(p) - np.pi / 2) < 1e-3:       # gimbal lock: roll := 0 (Pinocchio convention)
        r = 0.0
        y = -np.arctan2(R[0, 1], R[1, 1])
    else:
        y = np.arctan2(R[1, 0], R[0, 0])
        r = np.arctan2(R[2, 1], R[2, 2])
    return np.array([r, p, y])


def quat_to_matrix(xyzw: np.ndarray) -> np.ndarray:
    x, y, z, w = np.asarray(xyzw, dtype=float) / np.linalg.norm(xyzw)
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
    ])


def matrix_to_quat(R: np.ndarray) -> np.ndarray:
    """Rotation matrix -> (x, y, z, w), w >= 0 branch-stable (Shepperd)."""
    tr = np.trace(R)
    if tr > 0:
        s = 2.0 * np.sqrt(1.0 + tr)
        q = np.array([(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = 2.0 * np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k])
        q = np.zeros(4)
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    return q / np.linalg.norm(q)


def quat_to_ypr_state(q_full: np.ndarray) -> np.ndarray:
    """[p, quat(xyzw), joints] -> [p, yaw, pitch, roll, joints]  (transform.py:4-15)."""
    rpy = matrix_to_rpy(quat_to_matrix(q_full[3:7]))
    return np.hstack([q_full[:3], rpy[::-1], q_full[7:]])


def ypr_to_quat_state(q_euler: np.ndarray) -> np.ndarray:
    """[p, yaw, pitch, roll, joints] -> [p, quat(xyzw), joints]  (transform.py:17-27)."""
    R = rpy_to_matrix(q_euler[3:6][::-1])
    return np.hstack([q_euler[:3], matrix_to_quat(R), q_euler[6:]])


def ypr_to_quat_state_batched(q_euler_batch: np.ndarray) -> np.ndarray:
    return np.stack([ypr_to_quat_state(q) for q in q_euler_batch])


def _euler_rate_map(ypr: np.ndarray) -> np.ndarray:
    """E with w_local = E @ [yaw_dot, pitch_dot, roll_dot]  (transform.py:80-86)."""
    cx, sx, cy, sy = np.cos(ypr[2]), np.sin(ypr[2]), np.cos(ypr[1]), np.sin(ypr[1])
    return np.array([[-sy, 0., 1.], [cy * sx, cx, 0.], [cx * cy, -sx, 0.]])


def euler_derivative_to_local_angular(ypr_euler: np.ndarray, v_euler: np.ndarray) -> np.ndarray:
    return _euler_rate_map(ypr_euler) @ v_euler


def local_angular_to_euler_derivative(ypr_euler: np.ndarray, w_local: np.ndarray) -> np.ndarray:
    """Inverse map (transform.py:72-78); singular at pitch = +-pi/2 like the reference."""
    cx, sx, cy, sy = np.cos(ypr_euler[2]), np.sin(ypr_euler[2]), np.cos(ypr_euler[1]), np.sin(ypr_euler[1])
    T = np.array([[0., sx / cy, cx / cy], [0., cx, -sx], [1., sx * sy / cy, cx * sy / cy]])
    return T @ w_local


def v_global_linear_to_local_linear(q_euler: np.ndarray, v_glob_linear: np.ndarray) -> np.ndarray:
    """Rotates entries 3:6 into the base frame (transform.py:46-55, as written upstream)."""
    out = v_glob_linear.copy()
    out[3:6] = rpy_to_matrix(q_euler[3:6][::-1]).T @ v_glob_linear[3:6]
    return out


def v_local_linear_to_global_linear_batched(q_euler: np.ndarray, v_local_linear: np.ndarray) -> np.ndarray:
    out = v_local_linear.copy()
    for i, (ypr, w) in enumerate(zip(q_euler[:, 3:6], v_local_linear[:, 3:6])):
        out[i, 3:6] = rpy_to_matrix(ypr[::-1]) @ w
    return out


def v_to_euler_derivative(q_euler: np.ndarray, v: np.ndarray) -> np.ndarray:
    return np.concatenate((
        rpy_to_matrix(q_euler[3:6][::-1]) @ v[:3],
        local_angular_to_euler_derivative(q_euler[3:6], v[3:6]).flatten(),
        v[6:],
    ))
