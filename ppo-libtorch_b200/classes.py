"""The reference's model / distribution class surface over the C ABI (Python host side; the C++ facade is host/):

    Categorical(logits)                Distributions/Categorical.h:8-29   sample / log_prob / entropy / mode / mean / variance /
    CategoricalMasked(logits, masks)   Distributions/CategoricalMasked.h  enumerate_support, public m_logits / m_probs / m_num_events
    Agent(obs_size, action_size)       PPO/Agent.h:30-54                  getValue / getActionAndValueDiscrete / getActionAndValueMasked
    AgentOutput(action, logprob, entropy, value)   PPO/Agent.h:22-27

Same names, argument meaning and return shapes (action int64 [N] or [N, heads], logprob / entropy [N], value [N, 1]); every
number comes from one sm_100a kernel launch (ppo_categorical / ppo_policy_forward / ppo_value_forward) - there is no
torch arithmetic here and no CPU fallback.  Sampling draws from the library's counter-based Philox stream (seed, row,
call counter) or from injected uniforms (LibTorch's multinomial stream is not reproducible)."""
import collections
import math

import torch

from . import capi, ops
from .trainer import orthogonal_init_params

AgentOutput = collections.namedtuple("AgentOutput", ["action", "logprob", "entropy", "value"])


class Categorical:
    """Categorical over the last dimension of `logits` [N, K] (Distributions/Categorical.cpp:28-39).  entropy() keeps the
    reference's clamp quirk (== -FLT_MIN per row, :114-117) unless exact_entropy=True."""
    _masked = False

    def __init__(self, logits, device=None, seed=0, exact_entropy=False):
        if not (torch.is_tensor(logits) and logits.is_cuda):
            raise capi.PpoError("Categorical needs CUDA logits: the B200 path has no CPU fallback")
        self._x = logits.detach().to(torch.float32).reshape(-1, logits.shape[-1]).contiguous()
        self._mask = None
        self.m_num_events = self._x.shape[-1]
        self._seed, self._calls = seed, 0
        self._emode = capi.ENTROPY_EXACT if exact_entropy else capi.ENTROPY_REFERENCE
        self._cache = None
        self._norm = None

    def _eval(self, actions=None, uniforms=None):
        fresh = actions is None
        if not fresh:
            actions = actions.to(torch.int32).reshape(-1, 1).contiguous()
        if fresh:
            self._calls += 1
        return ops.categorical(self._x, [self.m_num_events], mask=self._mask, actions=actions, uniforms=uniforms,
                               seed=self._seed, counter=self._calls, entropy_mode=self._emode, want_probs=True)

    def _stats(self):
        if self._cache is None:   # any evaluation returns probs / entropy; actions = zeros is a valid choice of support
            zeros = torch.zeros(self._x.shape[0], dtype=torch.int32, device=self._x.device)
            self._cache = self._eval(actions=zeros)
        return self._cache

    @property
    def m_probs(self):            # softmax (Categorical.cpp:35)
        return self._stats()[3]

    @property
    def m_logits(self):           # logits - logsumexp (:33), from the kernel (ppo_categorical_normalize): finite wherever the
        if self._norm is None:    # reference's are (a masked logit reads -1e8 - lse, not log(0))
            self._norm = ops.categorical_normalize(self._x, [self.m_num_events], mask=self._mask)[0]
        return self._norm

    def sample(self, uniforms=None):
        """int64 [N]; `uniforms` [N] in [0,1) injects the draws (parity tests)"""
        u = None if uniforms is None else uniforms.to(torch.float32).reshape(-1, 1).contiguous()
        return self._eval(uniforms=u)[0].reshape(-1).to(torch.int64)

    def log_prob(self, value):    # gather of the normalised logits (:92-101)
        return self._eval(actions=value)[1]

    def entropy(self):            # :112-119 / CategoricalMasked.cpp:127-144
        return self._stats()[2]

    def mode(self):               # argmax of probs
        return torch.argmax(self.m_probs, dim=-1)

    def mean(self):               # undefined for a categorical: NaN like the reference
        return torch.full((self._x.shape[0],), math.nan, device=self._x.device)

    variance = mean

    def enumerate_support(self):
        return torch.arange(self.m_num_events, device=self._x.device).view(-1, 1).expand(-1, self._x.shape[0])

    @staticmethod
    def logits_to_probs(logits):
        return Categorical(logits).m_probs


class CategoricalMasked(Categorical):
    """where(mask, logits, -1e8) then Categorical; entropy sums valid terms only (CategoricalMasked.cpp:31-46, :127-144)"""
    _masked = True

    def __init__(self, logits, masks, device=None, seed=0):
        super().__init__(logits, device, seed, exact_entropy=False)
        self.m_masks = masks.to(torch.bool)
        self._mask = self.m_masks.reshape(-1, self.m_num_events).to(torch.uint8).contiguous()


class Agent:
    """Two independent 3-layer tanh MLPs (PPO/Agent.cpp:25-59), orthogonal init (:91-99); parameters are ONE flat fp32
    slab in the reference's registration order (the checkpoint layout).  action_size is an int (one head) or an nvec."""

    def __init__(self, obs_size, action_size, device=None, hidden=64, seed=1):
        if not torch.cuda.is_available():
            raise capi.PpoError("Agent needs a CUDA device: the B200 path has no CPU fallback")
        self.obs_size, self.hidden = int(obs_size), int(hidden)
        self.m_actionSpace = [int(action_size)] if isinstance(action_size, int) else [int(k) for k in action_size]
        self.params = torch.from_numpy(orthogonal_init_params(self.obs_size, self.hidden, self.m_actionSpace, seed)).cuda()
        self._seed, self._calls = seed, 0

    def parameters(self):
        """12 views of the slab, reference order and shapes (m_Critic.* then m_Actor.*)"""
        O, H, K = self.obs_size, self.hidden, sum(self.m_actionSpace)
        shapes = [(H, O), (H,), (H, H), (H,), (1, H), (1,), (H, O), (H,), (H, H), (H,), (K, H), (K,)]
        out, off = [], 0
        for s in shapes:
            n = math.prod(s)
            out.append(self.params[off:off + n].view(*s))
            off += n
        return out

    def getValue(self, x):
        return ops.value_forward(self.params, self.obs_size, self.hidden, self.m_actionSpace, x.contiguous()).view(-1, 1)

    def _forward(self, x, mask, action, uniforms):
        if action is not None:
            action = action.to(torch.int32).reshape(x.shape[0], -1).contiguous()
        else:
            self._calls += 1
        if mask is not None:
            mask = mask.to(torch.uint8).reshape(x.shape[0], -1).contiguous()
        a, lp, ent, val = ops.policy_forward(self.params, self.obs_size, self.hidden, self.m_actionSpace, x.contiguous(),
                                             mask=mask, actions=action, uniforms=uniforms, seed=self._seed,
                                             counter=self._calls)
        return a.to(torch.int64), lp, ent, val.view(-1, 1)

    def getActionAndValueDiscrete(self, x, action=None, uniforms=None):
        a, lp, ent, val = self._forward(x, None, action, uniforms)
        return AgentOutput(a.reshape(-1), lp, ent, val)

    def getActionAndValueMasked(self, x, action_masks, action=None, uniforms=None):
        a, lp, ent, val = self._forward(x, action_masks, action, uniforms)
        return AgentOutput(a, lp, ent, val)          # action [N, heads] like action.t() (Agent.cpp:168)
