"""Network specs standing in for dqn_zoo/networks.py on the learner hot path.

The reference builds haiku modules (networks.py:181-221,338-363); the B200 path needs only the
architecture parameters, so `dqn_atari_network` / `double_dqn_atari_network` return a light
`NatureCnnSpec` that the agent hands to the CUDA learner.  `init_params` reproduces the
reference initialisation scheme (networks.py:58-79): every weight and bias ~ U(-c, c) with
c = 1/sqrt(fan_in).  (The reference draws them from a jax PRNGKey; here a NumPy Generator.)
"""
import collections
import typing

import numpy as np

from dqn_mgsc_zoo_b200.learner import param_shapes

NATURE_HAIKU_NAMES = collections.OrderedDict([
    ("conv1", "sequential/sequential/conv2_d"), ("conv2", "sequential/sequential/conv2_d_1"),
    ("conv3", "sequential/sequential/conv2_d_2"), ("fc1", "sequential/sequential_1/linear"),
    ("fc2", "sequential/sequential_1/linear_1")])
_FAN_IN = {"conv1": 8 * 8 * 4, "conv2": 4 * 4 * 32, "conv3": 3 * 3 * 64, "fc1": 3136, "fc2": 512}


class NatureCnnSpec(typing.NamedTuple):
    num_actions: int
    shared_bias: bool = False


class RmsPropSpec(typing.NamedTuple):
    """Stands in for optax.rmsprop(learning_rate, decay, eps, centered=True)
    (dqn/run_atari.py:208-213); the update itself is the fused CUDA kernel."""
    learning_rate: float = 0.00025
    decay: float = 0.95
    eps: float = 0.01 / 32 ** 2
    centered: bool = True


class AdamSpec(typing.NamedTuple):
    """Stands in for optax.adam(learning_rate) (dqn_mgsc_batched/run_atari.py:241-243)."""
    learning_rate: float = 0.00025
    b1: float = 0.9
    b2: float = 0.999
    eps: float = 1e-8


def rmsprop(learning_rate, decay=0.9, eps=1e-8, centered=False) -> RmsPropSpec:
    """optax.rmsprop signature."""
    return RmsPropSpec(learning_rate, decay, eps, centered)


def adam(learning_rate, b1=0.9, b2=0.999, eps=1e-8) -> AdamSpec:
    """optax.adam signature."""
    return AdamSpec(learning_rate, b1, b2, eps)


def key_to_seed(rng_key):
    """A jax PRNGKey is a (2,) uint32 array; accept that, an int, or a SeedSequence-able list."""
    if isinstance(rng_key, (int, np.integer)):
        return int(rng_key)
    return [int(x) for x in np.asarray(rng_key).ravel()]


def dqn_atari_network(num_actions: int) -> NatureCnnSpec:
    """networks.py:352-363."""
    return NatureCnnSpec(num_actions, False)


def double_dqn_atari_network(num_actions: int) -> NatureCnnSpec:
    """networks.py:338-349 (single shared bias on the output layer)."""
    return NatureCnnSpec(num_actions, True)


def init_params(rng: np.random.Generator, num_actions: int, shared_bias: bool = False):
    params = collections.OrderedDict()
    for key, shape in param_shapes(num_actions, shared_bias).items():
        c = np.sqrt(1.0 / _FAN_IN[key.split(".")[0]])
        params[key] = rng.uniform(-c, c, size=shape).astype(np.float32)
    return params


def to_haiku_tree(params):
    """flat leaf dict -> {'sequential/.../conv2_d': {'w':..., 'b':...}, ...} (get_state format)."""
    tree = collections.OrderedDict()
    for key, value in params.items():
        layer, leaf = key.split(".")
        tree.setdefault(NATURE_HAIKU_NAMES[layer], {})[leaf] = np.asarray(value)
    return tree


def from_haiku_tree(tree):
    inv = {v: k for k, v in NATURE_HAIKU_NAMES.items()}
    flat = {}
    for module, leaves in tree.items():
        for leaf, value in leaves.items():
            flat[inv[module] + "." + leaf] = np.asarray(value)
    order = [l + "." + p for l in NATURE_HAIKU_NAMES for p in ("w", "b")]
    return collections.OrderedDict((k, flat[k]) for k in order)
