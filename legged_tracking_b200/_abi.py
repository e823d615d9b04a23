"""ctypes binding of include/lt_b200.h (the C ABI of liblt_b200.so).

The ctypes structures are generated from the header text itself, so the Python mirror cannot drift from
the C declaration: member order, array extents and pointer/scalar kinds all come from the one file.
This is the binding a maintainer of the reference would add to
go1_gym/envs/base/legged_robot_trajectory_tracking.py.

There is no CPU fallback: `load()` raises if the library is missing, and lt_create fails without an
sm_100 device.
"""
import ctypes as C
import os
import re

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(PKG_DIR)
HEADER = os.path.join(REPO, "include", "lt_b200.h")
POLICY_HEADER = os.path.join(REPO, "include", "lt_policy.h")       # actor-critic / PPO half of the same library
LIB_PATH = os.environ.get("LT_B200_LIB") or os.path.join(PKG_DIR, "liblt_b200.so")     # override: tuning variants only

_SCALARS = {"int32_t": C.c_int32, "uint32_t": C.c_uint32, "float": C.c_float, "int64_t": C.c_int64,
            "uint8_t": C.c_uint8, "double": C.c_double}


def _strip_comments(src):
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def parse_header(path=HEADER):
    """-> (defines: {name: int}, structs: {name: [(member, ctype)]}, member kinds for pointer structs)"""
    src = _strip_comments(open(path).read())
    if path == HEADER:
        src += "\n" + _strip_comments(open(POLICY_HEADER).read())
    defines = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(LT_\w+)\s+(-?\d+)\s", src)}
    enums = {}
    for body in re.findall(r"enum\s*\{(.*?)\}", src, flags=re.S) + re.findall(r"typedef enum\s*\{(.*?)\}", src, flags=re.S):
        val = -1
        for item in body.split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                name, v = [s.strip() for s in item.split("=")]
                val = int(v, 0)
            else:
                name, val = item, val + 1
            enums[name] = val
    structs, ptr_types, classes = {}, {}, {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\}\s*\1;", src, flags=re.S):
        name, body = m.group(1), m.group(2)
        fields, kinds = [], {}
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            const = decl.startswith("const ")
            if const:
                decl = decl[6:]
            typ, rest = decl.split(" ", 1)
            all_ptr = typ.endswith("*")
            typ = typ.rstrip("*")
            for item in rest.split(","):
                item = item.strip()
                is_ptr = all_ptr or item.startswith("*")
                item = item.lstrip("* ")
                am = re.match(r"(\w+)\[(\w+)\]$", item)
                if is_ptr:
                    fields.append((item, C.c_void_p))
                    kinds[item] = typ
                elif typ in classes:                      # a struct declared earlier in the header, embedded by value
                    fields.append((item, classes[typ]))
                elif am:
                    ext = am.group(2)
                    n = int(ext) if ext.isdigit() else defines[ext]
                    fields.append((am.group(1), _SCALARS[typ] * n))
                else:
                    fields.append((item, _SCALARS[typ]))
        structs[name] = fields
        ptr_types[name] = kinds
        classes[name] = type(name, (C.Structure,), {"_fields_": fields})
    return defines, enums, structs, ptr_types, classes


DEFINES, ENUMS, _STRUCTS, PTR_TYPES, _CLASSES = parse_header()


def _mk(name):
    return _CLASSES[name]


lt_config = _mk("lt_config")
lt_tables = _mk("lt_tables")
lt_sim_ptrs = _mk("lt_sim_ptrs")
lt_terrain_ptrs = _mk("lt_terrain_ptrs")
lt_state_ptrs = _mk("lt_state_ptrs")
lt_step_args = _mk("lt_step_args")
lt_rollout_ptrs = _mk("lt_rollout_ptrs")
lt_transition_ptrs = _mk("lt_transition_ptrs")
lt_plan_args = _mk("lt_plan_args")
lt_gait_ptrs = _mk("lt_gait_ptrs")
lt_corl_cfg = _mk("lt_corl_cfg")
lt_corl_ptrs = _mk("lt_corl_ptrs")
lt_cmd_cfg = _mk("lt_cmd_cfg")
lt_cmd_ptrs = _mk("lt_cmd_ptrs")
lt_vel_cfg = _mk("lt_vel_cfg")
lt_vel_ptrs = _mk("lt_vel_ptrs")
lt_vel_dev_cfg = _mk("lt_vel_dev_cfg")
lt_vel_dev_ptrs = _mk("lt_vel_dev_ptrs")
lt_policy_config = _mk("lt_policy_config")
lt_ppo_hparams = _mk("lt_ppo_hparams")
lt_ppo_batch_ptrs = _mk("lt_ppo_batch_ptrs")

EXPORTS = ["lt_abi_version", "lt_config_bytes", "lt_num_uniform_slots", "lt_create", "lt_destroy", "lt_last_error",
           "lt_bind_sim", "lt_bind_terrain", "lt_bind_state", "lt_compute_torques", "lt_post_physics_step", "lt_reset_idx",
           "lt_reward_sign_prepass", "lt_term_sum_buffer", "lt_set_term_scales", "lt_counters", "lt_rollout_store", "lt_gae",
           "lt_normalize_advantages", "lt_build_tunnel_tiles", "lt_plan_target_pose", "lt_measure_heights", "lt_step_contact_targets", "lt_corl_rewards", "lt_resample_commands", "lt_vel_pre", "lt_vel_reward", "lt_vel_observe", "lt_vel_measure_heights", "lt_vel_indexed_paths", "lt_vel_struct_bytes",
           # include/lt_policy.h
           "lt_policy_config_bytes", "lt_policy_create", "lt_policy_destroy", "lt_policy_last_error", "lt_policy_num_params",
           "lt_policy_bind_params", "lt_policy_refresh", "lt_policy_forward", "lt_policy_act", "lt_ppo_minibatch_grad",
           "lt_ppo_minibatch_apply", "lt_adaptation_grad", "lt_adaptation_apply", "lt_policy_adaptation_slice", "lt_policy_grad_scales", "lt_linear_f32", "lt_linear_tn_f32",
           "lt_linear_f32_static"]


def declared_functions(path=None):
    """Names of every function the headers declare (used by the symbol-export test)."""
    src = "\n".join(_strip_comments(open(h).read()) for h in ((HEADER, POLICY_HEADER) if path is None else (path,)))
    src = re.sub(r"typedef struct \w+ \{.*?\}\s*\w+;", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lt_\w+)\s*\(", src)))


_lib = None


class LtError(RuntimeError):
    pass


def load(path=LIB_PATH):
    """Load liblt_b200.so and declare its prototypes.  Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise LtError(f"{path} not found: build it with `python -m legged_tracking_b200.build` "
                      f"(nvcc, sm_100a).  There is no CPU fallback.")
    lib = C.CDLL(path)
    vp, i32 = C.c_void_p, C.c_int32
    lib.lt_abi_version.restype = C.c_int
    lib.lt_config_bytes.restype = C.c_int
    lib.lt_num_uniform_slots.argtypes = [C.POINTER(lt_config)]
    lib.lt_create.argtypes = [C.POINTER(lt_config), C.POINTER(lt_tables), C.POINTER(vp)]
    lib.lt_destroy.argtypes = [vp]
    lib.lt_destroy.restype = None
    lib.lt_last_error.argtypes = [vp]
    lib.lt_last_error.restype = C.c_char_p
    lib.lt_bind_sim.argtypes = [vp, C.POINTER(lt_sim_ptrs)]
    lib.lt_bind_terrain.argtypes = [vp, C.POINTER(lt_terrain_ptrs), i32]
    lib.lt_bind_state.argtypes = [vp, C.POINTER(lt_state_ptrs)]
    lib.lt_compute_torques.argtypes = [vp, vp, i32, vp]
    lib.lt_post_physics_step.argtypes = [vp, C.POINTER(lt_step_args), vp]
    lib.lt_reward_sign_prepass.argtypes = [vp, C.POINTER(lt_step_args), vp]
    lib.lt_term_sum_buffer.argtypes = [vp]
    lib.lt_term_sum_buffer.restype = vp
    lib.lt_reset_idx.argtypes = [vp, vp, i32, vp, vp, vp]
    lib.lt_set_term_scales.argtypes = [vp, C.POINTER(C.c_float), i32]
    lib.lt_counters.argtypes = [vp]
    lib.lt_counters.restype = vp
    lib.lt_rollout_store.argtypes = [C.POINTER(lt_rollout_ptrs), C.POINTER(lt_transition_ptrs), i32, i32, i32, i32, i32, i32,
                                     C.c_float, vp]
    lib.lt_gae.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, C.c_float, C.c_float, vp]
    lib.lt_normalize_advantages.argtypes = [vp, vp, C.c_int64, vp]
    lib.lt_build_tunnel_tiles.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    i64, f32 = C.c_int64, C.c_float
    lib.lt_measure_heights.argtypes = [vp, vp, vp]
    lib.lt_step_contact_targets.argtypes = [C.POINTER(lt_gait_ptrs), i32, i32, C.c_float, C.c_float, i32, vp]
    lib.lt_corl_rewards.argtypes = [C.POINTER(lt_corl_cfg), C.POINTER(lt_corl_ptrs), vp]
    lib.lt_resample_commands.argtypes = [C.POINTER(lt_cmd_cfg), C.POINTER(lt_cmd_ptrs), i32, vp]
    lib.lt_vel_pre.argtypes = [C.POINTER(lt_vel_cfg), C.POINTER(lt_vel_ptrs), vp]
    lib.lt_vel_indexed_paths.argtypes = [C.POINTER(lt_vel_cfg), C.POINTER(lt_vel_ptrs), C.POINTER(lt_vel_dev_cfg), C.POINTER(lt_vel_dev_ptrs), i32, vp,
                                         C.c_uint32, vp]
    lib.lt_vel_measure_heights.argtypes = [vp, vp, vp, i32, i32, i32, i32, f32, f32, f32, vp, vp, vp]
    lib.lt_vel_reward.argtypes = [C.POINTER(lt_vel_cfg), C.POINTER(lt_vel_ptrs), vp, vp]
    lib.lt_vel_observe.argtypes = [C.POINTER(lt_vel_cfg), C.POINTER(lt_vel_ptrs), vp, C.c_uint32, vp]
    lib.lt_plan_target_pose.argtypes = [C.POINTER(lt_plan_args), i32, i32, i32, i32, i32, C.POINTER(C.c_float), vp]
    lib.lt_policy_config_bytes.restype = C.c_int
    lib.lt_policy_create.argtypes = [C.POINTER(lt_policy_config), C.POINTER(vp)]
    lib.lt_policy_destroy.argtypes = [vp]
    lib.lt_policy_destroy.restype = None
    lib.lt_policy_last_error.argtypes = [vp]
    lib.lt_policy_last_error.restype = C.c_char_p
    lib.lt_policy_num_params.argtypes = [vp]
    lib.lt_policy_num_params.restype = i64
    lib.lt_policy_bind_params.argtypes = [vp, vp, vp]
    lib.lt_policy_refresh.argtypes = [vp, vp]
    lib.lt_policy_forward.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp]
    lib.lt_policy_act.argtypes = [vp, vp, vp, i32, vp, vp, C.c_uint64, C.c_uint64, vp, vp, vp, vp, vp, vp]
    lib.lt_ppo_minibatch_grad.argtypes = [vp, C.POINTER(lt_ppo_batch_ptrs), vp, i32, C.POINTER(lt_ppo_hparams), i32, vp, vp]
    lib.lt_ppo_minibatch_apply.argtypes = [vp, C.POINTER(lt_ppo_hparams), vp, vp, vp, vp, i32, vp, vp]
    lib.lt_adaptation_grad.argtypes = [vp, C.POINTER(lt_ppo_batch_ptrs), vp, i32, i32, i32, vp, vp]
    lib.lt_adaptation_apply.argtypes = [vp, C.POINTER(lt_ppo_hparams), f32, vp, vp, i32, vp]
    lib.lt_policy_adaptation_slice.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    lib.lt_policy_grad_scales.argtypes = [vp]
    lib.lt_policy_grad_scales.restype = vp
    lib.lt_linear_f32.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, vp]
    lib.lt_linear_tn_f32.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp]
    lib.lt_linear_f32_static.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp]
    lib.lt_vel_struct_bytes.argtypes = [i32]
    for which, cls in enumerate((lt_vel_cfg, lt_vel_ptrs, lt_vel_dev_cfg, lt_vel_dev_ptrs, lt_cmd_cfg)):
        if lib.lt_vel_struct_bytes(which) != C.sizeof(cls):
            raise LtError(f"sizeof({cls.__name__}) is {lib.lt_vel_struct_bytes(which)} in the library, {C.sizeof(cls)} in the binding: rebuild")
    if lib.lt_policy_config_bytes() != C.sizeof(lt_policy_config):
        raise LtError("sizeof(lt_policy_config) differs between the library and the binding: rebuild")
    if lib.lt_abi_version() != DEFINES["LT_ABI_VERSION"]:
        raise LtError(f"liblt_b200.so has ABI {lib.lt_abi_version()}, header has {DEFINES['LT_ABI_VERSION']}: rebuild")
    if lib.lt_config_bytes() != C.sizeof(lt_config):
        raise LtError(f"sizeof(lt_config) = {lib.lt_config_bytes()} in the library but {C.sizeof(lt_config)} in the binding")
    _lib = lib
    return lib


def check_policy(rc, handle=None, what="liblt_b200 policy"):
    if rc != 0:
        msg = load().lt_policy_last_error(handle)
        raise LtError(f"{what} failed with status {rc}: {msg.decode() if msg else '?'}")


def check(rc, handle=None, what="liblt_b200"):
    if rc != 0:
        msg = load().lt_last_error(handle)
        raise LtError(f"{what} failed with status {rc}: {msg.decode() if msg else '?'}")
