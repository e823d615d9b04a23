"""Fake `isaacgym.gymapi` (oracle shim, test infrastructure only).

A `FakeGym` singleton that answers the Isaac Gym calls made by
go1_gym/envs/base/legged_robot_trajectory_tracking.py and base_task.py.  The four
simulation tensors are plain torch CPU tensors with the layouts Isaac Gym
documents for acquire_actor_root_state_tensor / acquire_dof_state_tensor /
acquire_net_contact_force_tensor / acquire_rigid_body_state_tensor.

PhysX is replaced by a *replay source*: an object with
    on_simulate(gym)   -- called from gym.simulate(); writes the next substep's dof_state
    on_refresh_root(gym) -- called from refresh_actor_root_state_tensor(); writes root /
                            contact / rigid-body state for the finished policy step
Set it with `acquire_gym().set_replay(src)`.
"""
import numpy as np
import torch

SIM_PHYSX = 1
DOF_MODE_NONE = 0
FOLLOW_POSITION = 0
IMAGE_COLOR = 0
KEY_ESCAPE = 0
KEY_V = 1

# Go1 asset description (body / DOF order as Isaac Gym reports it for go1.urdf with
# collapse_fixed_joints=True; DOF order matches go1_gym_deploy/envs/lcm_traj_agent.py:67-71).
LEGS = ["FL", "FR", "RL", "RR"]
GO1_BODY_NAMES = ["base"] + [f"{l}_{p}" for l in LEGS for p in ("hip", "thigh", "calf", "foot")]
GO1_DOF_NAMES = [f"{l}_{p}_joint" for l in LEGS for p in ("hip", "thigh", "calf")]
# resources/robots/go1/urdf/go1.urdf:96,138,166
GO1_DOF_LOWER = [-0.802851455917, -1.0471975512, -2.69653369433] * 4
GO1_DOF_UPPER = [0.802851455917, 4.18879020479, -0.916297857297] * 4
GO1_DOF_VEL = [50.0, 28.0, 28.0] * 4
GO1_DOF_EFFORT = [33.5] * 12


class Vec3:
    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)


class Quat:
    def __init__(self, x=0.0, y=0.0, z=0.0, w=1.0):
        self.x, self.y, self.z, self.w = x, y, z, w

    @staticmethod
    def from_axis_angle(axis, angle):
        return Quat()


class Transform:
    def __init__(self, p=None, r=None):
        self.p = p if p is not None else Vec3()
        self.r = r if r is not None else Quat()


class _Bag:
    """Attribute bag used for option/param structs."""
    def __init__(self, **kw):
        self.__dict__.update(kw)


class AssetOptions(_Bag):
    pass


class CameraProperties(_Bag):
    pass


class PlaneParams(_Bag):
    pass


class TriangleMeshParams(_Bag):
    def __init__(self):
        super().__init__(transform=Transform())


class RigidShapeProperties(_Bag):
    def __init__(self):
        super().__init__(friction=1.0, restitution=0.0)


class RigidBodyProperties(_Bag):
    def __init__(self):
        super().__init__(mass=1.0, com=Vec3())


class SimType:
    pass


class PhysXParams(_Bag):
    pass


class SimParams:
    def __init__(self):
        # Isaac Gym stores dt as a C float: decimation * dt = 0.019999999552965164
        self.dt = float(np.float32(0.005))
        self.substeps = 1
        self.gravity = Vec3(0.0, 0.0, -9.81)
        self.up_axis = 1
        self.use_gpu_pipeline = True
        self.physx = PhysXParams()


class _Asset:
    def __init__(self, kind):
        self.kind = kind


class FakeGym:
    def __init__(self):
        self.replay = None
        self.num_envs_created = 0
        self.actors_per_env = {}
        self.sim_params = None
        self.tensors = {}
        self.calls = []          # log of write-back calls (set_*_indexed), for tests

    def set_replay(self, src):
        self.replay = src

    # ---- sim / assets -------------------------------------------------------
    def create_sim(self, dev, gdev, engine, sim_params):
        self.sim_params = sim_params
        self.num_envs_created = 0
        self.actors_per_env = {}
        self.tensors = {}
        return "sim"

    def prepare_sim(self, sim):
        pass

    def get_sim_params(self, sim):
        return self.sim_params

    def set_sim_params(self, sim, p):
        self.sim_params = p

    def add_ground(self, sim, params):
        pass

    def add_triangle_mesh(self, *a, **k):
        pass

    def load_asset(self, sim, root, file, options):
        return _Asset("arrow" if "arrow" in file else "robot")

    def create_box(self, *a, **k):
        return _Asset("box")

    def get_asset_dof_count(self, asset):
        return len(GO1_DOF_NAMES)

    def get_asset_rigid_body_count(self, asset):
        return len(GO1_BODY_NAMES)

    def get_asset_dof_names(self, asset):
        return list(GO1_DOF_NAMES)

    def get_asset_rigid_body_names(self, asset):
        return list(GO1_BODY_NAMES)

    def get_asset_dof_properties(self, asset):
        n = len(GO1_DOF_NAMES)
        props = np.zeros(n, dtype=[("lower", "f4"), ("upper", "f4"), ("velocity", "f4"), ("effort", "f4")])
        props["lower"] = GO1_DOF_LOWER
        props["upper"] = GO1_DOF_UPPER
        props["velocity"] = GO1_DOF_VEL
        props["effort"] = GO1_DOF_EFFORT
        return props

    def get_asset_rigid_shape_properties(self, asset):
        return [RigidShapeProperties() for _ in GO1_BODY_NAMES]

    def set_asset_rigid_shape_properties(self, asset, props):
        pass

    # ---- envs / actors ------------------------------------------------------
    def create_env(self, sim, lo, hi, per_row):
        h = self.num_envs_created
        self.num_envs_created += 1
        self.actors_per_env[h] = 0
        return h

    def create_actor(self, env, asset, pose, name, group, filt, seg=0):
        h = self.actors_per_env[env]
        self.actors_per_env[env] += 1
        return h

    def set_actor_dof_properties(self, *a, **k):
        pass

    def get_actor_rigid_body_properties(self, env, actor):
        return [RigidBodyProperties() for _ in GO1_BODY_NAMES]

    def set_actor_rigid_body_properties(self, *a, **k):
        pass

    def get_actor_rigid_shape_properties(self, env, actor):
        return [RigidShapeProperties() for _ in GO1_BODY_NAMES]

    def set_actor_rigid_shape_properties(self, *a, **k):
        pass

    def find_actor_rigid_body_handle(self, env, actor, name):
        return GO1_BODY_NAMES.index(name)

    # ---- state tensors ------------------------------------------------------
    def _alloc(self):
        if self.tensors:
            return
        n = self.num_envs_created
        a = self.actors_per_env[0]
        nb = len(GO1_BODY_NAMES) + (a - 1)          # robot bodies + one body per extra actor (arrow)
        root = torch.zeros(n * a, 13)
        root[:, 6] = 1.0
        self.tensors = dict(
            root=root,
            dof=torch.zeros(n * len(GO1_DOF_NAMES), 2),
            contact=torch.zeros(n * nb, 3),
            rigid=torch.zeros(n * nb, 13),
        )
        self.num_bodies_per_env = nb
        self.num_actors_per_env = a

    def acquire_actor_root_state_tensor(self, sim):
        self._alloc()
        return self.tensors["root"]

    def acquire_dof_state_tensor(self, sim):
        self._alloc()
        return self.tensors["dof"]

    def acquire_net_contact_force_tensor(self, sim):
        self._alloc()
        return self.tensors["contact"]

    def acquire_rigid_body_state_tensor(self, sim):
        self._alloc()
        return self.tensors["rigid"]

    def refresh_dof_state_tensor(self, sim):
        pass

    def refresh_actor_root_state_tensor(self, sim):
        if self.replay is not None:
            self.replay.on_refresh_root(self)

    def refresh_net_contact_force_tensor(self, sim):
        pass

    def refresh_rigid_body_state_tensor(self, sim):
        pass

    def simulate(self, sim):
        if self.replay is not None:
            self.replay.on_simulate(self)

    def fetch_results(self, sim, wait):
        pass

    def set_dof_actuation_force_tensor(self, sim, t):
        self.last_torques = t

    def set_dof_state_tensor_indexed(self, sim, t, ids, n):
        self.calls.append(("dof_indexed", ids.clone()))

    def set_actor_root_state_tensor_indexed(self, sim, t, ids, n):
        self.calls.append(("root_indexed", ids.clone()))

    def set_actor_root_state_tensor(self, sim, t):
        self.calls.append(("root_all", None))

    # ---- graphics: no-ops ---------------------------------------------------
    def render_all_camera_sensors(self, sim):
        pass

    def step_graphics(self, sim):
        pass

    def create_camera_sensor(self, *a, **k):
        return 0

    def set_camera_location(self, *a, **k):
        pass

    def attach_camera_to_body(self, *a, **k):
        pass

    def get_camera_image(self, *a, **k):
        return np.zeros((240, 360 * 4), dtype=np.uint8)

    def create_viewer(self, *a, **k):
        return None

    def destroy_viewer(self, *a, **k):
        pass

    def destroy_sim(self, *a, **k):
        pass

    def clear_lines(self, *a, **k):
        pass


_GYM = None


def acquire_gym():
    global _GYM
    if _GYM is None:
        _GYM = FakeGym()
    return _GYM
