"""Configuration surface of the reference that the env path consumes.

``TrainConfig`` mirrors the fields of the reference's ``classes/train_config.py:6-68`` that reach the
environment constructor (``panda_tasks.py:142-159``); learner / wandb fields are kept as plain data so that a
reference config object can be passed through unchanged (any object with these attributes works).
"""
from dataclasses import dataclass, field
from typing import Optional


@dataclass
class TrainConfig:
    name: str = "default"
    job_type: str = "train"
    group: str = "default"
    algorithm: str = "TQC"
    policy_type: str = "MultiInputPolicy"
    learning_starts: int = 10000
    prior_steps: int = 0
    seed: int = 0
    n_envs: int = 8
    env_name: str = "PandaReachAO-v3"
    randomize_robot_pose: bool = False
    truncate_on_collision: bool = True
    terminate_on_success: bool = True
    fixed_target: Optional[list] = None
    reward_type: str = "sparse"
    collision_reward: int = -100
    goal_condition: str = "reach"
    ee_error_thresholds: list = field(default_factory=lambda: [0.05, 0.05, 0.05])
    speed_thresholds: list = field(default_factory=lambda: [0.5, 0.1, 0.01])
    safety_distance: float = 0.0
    max_timesteps: int = 600_000
    max_ep_steps: list = field(default_factory=lambda: [50, 75, 100])
    n_substeps: int = 20
    stages: list = field(default_factory=lambda: ["reachao1", "reachao2", "reachao3"])
    success_thresholds: list = field(default_factory=lambda: [0.9, 0.9, 1.0])
    eval_freq: int = 10_000
    n_eval_episodes: int = 100
    obs_type: tuple = ("ee", "js")
    control_type: str = "js"
    action_limiter: str = "clip"
    limiter: str = "sim"
    task_observations: dict = field(default_factory=lambda: {"obstacles": "vectors+closest_per_link", "prior": None})
    render: bool = False
    show_goal_space: bool = False
    show_debug_labels: bool = False
    debug_collision: bool = False


CONTROL = {"js": 0, "joints": 0, "ee": 1, "jsd": 2, "pcc": 3}
REWARD_AO = {"sparse": 0, "wang": 1, "kumar_her": 2, "kumar_optim": 3, "kumar": 4}
REWARD_CLASSIC = {"sparse": 0, "dense": 1}
OBSTACLE_OBS = {"vectors+closest_per_link": 1, "vectors": 2, "closest_per_link": 3, "closest": 4, "vectors+past": 5}
