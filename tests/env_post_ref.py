"""numpy restatement of the two GPUDriveTorchEnv post-processing steps the engine also offers on the device
(gd_shaped_reward, gd_expert_actions): gpudrive/env/env_torch.py:469-604 (get_rewards) and :1445-1509
(get_expert_actions) with LogTrajectory (gpudrive/datatypes/trajectory.py:21-39). Test infrastructure only."""
import numpy as np

TRAJ_LEN = 91
f32 = np.float32


def shaped_reward(info, reward, traj, abs_obs, mode, collision_weight=-0.5, goal_achieved_weight=1.0,
                  off_road_weight=-0.5, log_distance_weight=0.01, agent_weights=None, world_time_steps=None):
    """info i32 [W,S,5], reward f32 [W,S,1], traj f32 [W,S,1456], abs_obs f32 [W,S,14] -> f32 [W,S]."""
    if mode == "sparse_on_goal_achieved":
        return reward[:, :, 0].copy()
    off_road = info[:, :, 0].astype(f32)
    collided = info[:, :, 1].astype(f32) + info[:, :, 2].astype(f32)
    goal = info[:, :, 3].astype(f32)
    if mode == "reward_conditioned":
        cw, gw, ow = agent_weights[:, :, 0], agent_weights[:, :, 1], agent_weights[:, :, 2]
    else:
        cw, gw, ow = f32(collision_weight), f32(goal_achieved_weight), f32(off_road_weight)
    r = (cw * collided + gw * goal) + ow * off_road
    if mode == "distance_to_logs":
        W, S = info.shape[:2]
        pos = traj[:, :, :2 * TRAJ_LEN].reshape(W, S, TRAJ_LEN, 2)
        log_pos = np.stack([pos[w, :, int(world_time_steps[w]), :] for w in range(W)])
        d = log_pos - abs_obs[:, :, 0:2]
        dist = np.sqrt(d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1])
        r = r + f32(log_distance_weight) * np.exp(-dist)
    return r.astype(f32)


def expert_actions(traj, model):
    """traj f32 [W,S,1456]; model: 0 classic, 1 bicycle, 2 delta_local, 3 state -> f32 [W,S,91,3|10]."""
    W, S = traj.shape[:2]
    T = TRAJ_LEN
    if model == 3:
        pos = traj[:, :, :2 * T].reshape(W, S, T, 2)
        vel = traj[:, :, 2 * T:4 * T].reshape(W, S, T, 2)
        yaw = traj[:, :, 4 * T:5 * T].reshape(W, S, T, 1)
        return np.concatenate([pos, np.ones((W, S, T, 1), f32), yaw, vel, np.zeros((W, S, T, 4), f32)], axis=-1)
    a = traj[:, :, 6 * T:16 * T].reshape(W, S, T, 10)[..., :3].copy()
    a[..., 0] = np.clip(a[..., 0], -6, 6)
    if model == 2:
        a[..., 1] = np.clip(a[..., 1], -6, 6)
        a[..., 2] = np.clip(a[..., 2], -np.pi, np.pi).astype(f32)
    else:
        a[..., 1] = np.clip(a[..., 1], f32(-0.3), f32(0.3))
    return a
