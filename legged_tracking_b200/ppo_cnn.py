"""go1_gym_learn/ppo_cse_cnn behind the reference's names: the PPO variant scripts/train.py selects
when --old_ppo is absent (scripts/train.py:17-26).  CNN below = go1_gym_learn/ppo_cse_cnn/actor_critic.py.

Against ppo_cse it differs in ONE place: the three bodies read `process_obs_history(obs_history)` (CNN:181-198) instead of the
observation history -- the height columns of the last observation row go through HeightMapEncoder (CNN:27-62) first.  PPO,
RolloutStorage and Runner are the same code as ppo_cse's (the reference's two copies differ by the one `process_obs_history`
call in the adaptation step, ppo_cse_cnn/ppo.py:170-171, and by logging), so they are shared with ppo.py here.

Built: the default form -- AC_Args.use_cnn = False (HeightMapEncoder = Linear, act, Linear, act) and AC_Args.use_gru = False
(recurrent_latent_embedding = Identity).  All of it runs inside liblt_b200 (lt_policy_config.height_map_size > 0): the encoder is
two more tcgen05 contractions in front of the bodies, its gradient the sum of the three bodies' first-layer input gradients.
Not built: use_cnn = True (two 3x3 convolutions + max-pools, CNN:36-45) and use_gru = True (nn.GRU over the history rows,
CNN:100-101) -- both raise here; there is no torch fallback.  (In the reference the Conv2d form ends in Linear(3360, emb) and
fails for the 2 x 21 x 11 map scripts/train.py configures; the GRU is seq-first and fed [batch, H, features], so it recurs over
the batch dimension: DESIGN.md section 8.)
"""
import math

from . import ppo as _ppo
from .params import PrefixProto
from .ppo import PPO, PPO_Args, RolloutStorage, RunnerArgs          # noqa: F401  (same classes, re-exported under this package)


class AC_Args(PrefixProto, cli=False):          # CNN:8-24
    init_noise_std = 1.0
    actor_hidden_dims = [512, 256, 128]
    critic_hidden_dims = [512, 256, 128]
    activation = 'elu'
    adaptation_module_branch_hidden_dims = [256, 128]
    use_decoder = False
    use_cnn = False
    use_gru = False
    height_map_shape = (2, 61, 31)
    cnn_num_embedding = 256
    lstm_num_embedding = 256


class ActorCritic(_ppo.ActorCritic):
    """CNN:65-238.  state_dict keys as the reference's: 'std', 'height_map_encoder.model.{0,2}.{weight,bias}',
    'adaptation_module.*', 'actor_body.*', 'critic_body.*' (recurrent_latent_embedding is an Identity and holds none)."""
    ARGS = AC_Args
    ENCODER_HIDDEN = 256                        # CNN:47

    def _encoder_spec(self, args):
        if args.use_cnn:
            raise NotImplementedError("AC_Args.use_cnn = True (Conv2d / MaxPool2d encoder, CNN:36-45) has no kernel here")
        if args.use_gru:
            raise NotImplementedError("AC_Args.use_gru = True (nn.GRU over the history rows, CNN:100-101) has no kernel here")
        hm = int(math.prod(args.height_map_shape))
        if hm > self.num_obs:
            # the reference slices [-hm:] and then reshapes: with hm > num_obs that only works by accident of the batch size
            raise ValueError(f"height_map_shape {tuple(args.height_map_shape)} has {hm} cells but an observation row only {self.num_obs} columns")
        self.height_map_shape = tuple(int(v) for v in args.height_map_shape)
        self.cnn_num_embedding = int(args.cnn_num_embedding)
        self.lstm_input_dim = self.num_obs - hm + self.cnn_num_embedding           # CNN:88-90
        self.lstm_num_embedding = self.lstm_input_dim
        return dict(height_map_size=hm, hidden=self.ENCODER_HIDDEN, embedding=self.cnn_num_embedding)

    def export_module(self, name):
        if name == "height_map_encoder.model":
            # both Linear layers are followed by the activation (CNN:46-52)
            import torch.nn as nn
            seq = super().export_module(name)
            act = type(seq[1])
            return nn.Sequential(*list(seq), act())
        return super().export_module(name)


class Runner(_ppo.Runner):
    ACTOR_CRITIC = ActorCritic

    def __init__(self, env, device='cuda:0', runner_args=RunnerArgs, ac_args=AC_Args, **kw):
        super().__init__(env, device=device, runner_args=runner_args, ac_args=ac_args, **kw)
