"""Interface parity with dqn_zoo/replay_jax.py:789-994 (MGSCDistribution, MGSCFiFoTransitionReplay).

No agent in the reference imports that module; it differs from replay_circular only
in taking a JAX `rng_key` and sampling with jax.random.choice (threefry).  The classes here expose
the same method set on top of the device logit ring of replay_circular.py; the key seeds a NumPy
Generator, so the *index stream* is not the threefry stream of the reference (not promised).
"""
import numpy as np

from dqn_mgsc_zoo_b200 import replay_circular
from dqn_mgsc_zoo_b200.networks import key_to_seed
from dqn_mgsc_zoo_b200.replay import Transition  # noqa: F401


class MGSCDistribution:
    """replay_jax.py:789-885 method set over a device logit ring (relative indices)."""

    def __init__(self, rng_key, max_capacity: int, device=0, stream=0):
        self._ring = replay_circular.CircularLogitBuffer(
            max_capacity, np.random.default_rng(key_to_seed(rng_key)), device=device, stream=stream)

    def ensure_capacity(self, capacity: int) -> None:
        if capacity > self._ring.capacity:
            raise BufferError("capacity %d exceeds max_capacity %d" % (capacity, self._ring.capacity))

    def add_priorities(self, indices, priorities) -> None:
        for i, p in zip(indices, priorities):
            if i != self._ring.size:
                raise IndexError("priorities are appended: expected index %d, got %d" % (self._ring.size, i))
            self._ring.add(np.float32(p))

    def remove_priorities(self, indices) -> None:
        for i in indices:
            if i != 0:
                raise IndexError("only the oldest priority (index 0) can be removed, got %d" % i)
            self._ring.popleft(want_value=False)

    def update_priorities(self, indices, priorities) -> None:
        self._ring[np.asarray(indices)] = priorities

    def get_priorities(self, indices) -> np.ndarray:
        return self._ring[np.asarray(indices)]

    def probabilities(self) -> np.ndarray:
        return self._ring.as_probs()

    def sample(self, size: int) -> np.ndarray:
        return self._ring.sample(size)

    def uniform_sample(self, size: int, replace: bool = True) -> np.ndarray:
        return self._ring.sample_uniform(size, replace=replace)

    @property
    def _logits(self):
        return self._ring._logits

    @property
    def capacity(self) -> int:
        return self._ring.capacity

    @property
    def size(self) -> int:
        return self._ring.size

    def get_state(self):
        return self._ring.get_state()

    def set_state(self, state) -> None:
        self._ring.set_state(state)

    def check_valid(self):
        return True, ''


class MGSCFiFoTransitionReplay(replay_circular.MGSCFiFoTransitionReplay):
    """replay_jax.py:887-994: same methods as the circular replay, constructed from a PRNG key."""

    def __init__(self, capacity, structure, rng_key, encoder=None, decoder=None, device=0, stream=0,
                 initial_planes=0):
        super().__init__(capacity, structure, np.random.default_rng(key_to_seed(rng_key)), encoder, decoder,
                         device=device, stream=stream, initial_planes=initial_planes)

    def slice_of_ids_transitions_and_logits(self, size: int):
        raise NotImplementedError("TODO")   # replay_jax.py:952-957 raises the same
