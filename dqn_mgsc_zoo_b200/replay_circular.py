"""Drop-in for the MGSC replays of dqn_zoo/replay_circular.py, on the device.

Mirrors (same names, arguments, errors, Generator consumption):
  CircularLogitBuffer              replay_circular.py:148-248
  MGSCReservoirDistribution        replay_circular.py:500-560
  MGSCFiFoTransitionReplay         replay_circular.py:1217-1312
  MGSCReservoirTransitionReplay    replay_circular.py:568-664

The reference walks its whole float32 logit array on the host for every add (logsumexp) and
every sample (softmax, float64 cumsum, searchsorted): 1.94 ms / 57 ms at capacity 1e6.  Here the
array lives in HBM (`dqnb_logits_*`, csrc/logits.cu); ring heads, sizes and the shared NumPy
Generator stay on the host and are consumed draw-for-draw like the reference:
  weighted sample : Generator.choice(cap, n, p=p) draws Generator.random(n) once -> we draw the same
                    n doubles and search the device CDF with them;
  meta batch      : Generator.choice(size, M, replace=False) has no logits in it -> unchanged, host.
Frames go to the same HBM frame store as the uniform replays (ring slot = absolute index).
Indices equal the reference's unless a uniform falls within float rounding of a CDF boundary
(the device's expf/logf and summation order differ from NumPy's in the last ulp); logits agree
to float32 rounding.
"""
import ctypes
from typing import Optional

import numpy as np

from dqn_mgsc_zoo_b200 import _lib
from dqn_mgsc_zoo_b200 import replay as _uniform
from dqn_mgsc_zoo_b200.replay import (ItemStorage, Transition, TransitionAccumulator,  # noqa: F401  (re-exported:
                                      NStepTransitionAccumulator, compress_array,     # the MGSC runners import this
                                      uncompress_array)                               # module as `replay_lib`)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


class CircularBuffer:
    """Host ring of arbitrary items (replay_circular.py:90-146); the MGSC replays here keep frame items in
    the HBM store instead, this class is the module-surface twin for code that uses it directly."""

    def __init__(self, capacity: int):
        self._list = [None] * capacity
        self._capacity = capacity
        self._size = 0
        self._left_head = 0
        self._right_head = 0

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def size(self) -> int:
        return self._size

    def is_full(self) -> bool:
        return self._size == self._capacity

    def add(self, item) -> None:
        if self.is_full():
            raise BufferError("Buffer is full and cannot be added to. Pop an item first.")
        self._list[self._right_head] = item
        self._right_head = (self._right_head + 1) % self._capacity
        self._size += 1

    def popleft(self):
        if self._size == 0:
            raise BufferError("Buffer is empty and cannot be popped from. Add an item first.")
        item = self._list[self._left_head]
        self._left_head = (self._left_head + 1) % self._capacity
        self._size -= 1
        return item

    def __getitem__(self, key: int):
        if self._size == 0:
            raise BufferError("Buffer is empty and cannot be indexed. Add an item first.")
        return self._list[(self._left_head + key) % self._capacity]

    def __len__(self):
        return self._size

    def get_state(self):
        return {'list': self._list, 'capacity': self._capacity, 'size': self._size,
                'left_head': self._left_head, 'right_head': self._right_head}

    def set_state(self, state) -> None:
        self._list = state['list']
        self._capacity = state['capacity']
        self._size = state['size']
        self._left_head = state['left_head']
        self._right_head = state['right_head']


class UniformDistribution(_uniform.UniformDistribution):
    """replay_circular.py:251-324: the `np.random.Generator` twin (`integers` instead of `randint`)."""

    def _draw(self, high, size):
        return self._random_state.integers(high, size=size)          # replay_circular.py:285


class TransitionReplay(_uniform.TransitionReplay):
    """replay_circular.py:327-407 (FIFO, Generator)."""
    _DISTRIBUTION = UniformDistribution


class ReservoirTransitionReplay(_uniform.ReservoirTransitionReplay):
    """replay_circular.py:410-497 (Algorithm R, Generator)."""
    _DISTRIBUTION = UniformDistribution

    def _draw_replacement(self):
        return self._random_state.integers(0, self._t)                # replay_circular.py:436


class PrioritizedDistribution(_uniform.PrioritizedDistribution):
    """replay_circular.py:893-1116: the Generator twin (`integers` for the uniform positions)."""

    def _draw_positions(self, size):
        return self._random_state.integers(self.size, size=size)      # replay_circular.py:1017


class PrioritizedTransitionReplay(_uniform.PrioritizedTransitionReplay):
    """replay_circular.py:1315-1430."""
    _DISTRIBUTION = PrioritizedDistribution


SumTree = _uniform.SumTree
importance_sampling_weights = _uniform.importance_sampling_weights


class _DeviceLogits:
    """float32 logits[capacity] in device memory; absolute indices."""

    def __init__(self, capacity, device=0, stream=0):
        self.capacity, self.device, self.stream = int(capacity), device, stream
        h = ctypes.c_void_p()
        _lib.call("dqnb_logits_create", self.capacity, device, ctypes.byref(h))
        self._h = h

    def close(self):
        if self._h is not None:
            _lib.call("dqnb_logits_destroy", self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def fill(self, index, size):
        _lib.call("dqnb_logits_fill", self._h, int(index), int(size), self.stream)

    def set(self, indices, values):
        idx = _i64(np.atleast_1d(indices))
        val = np.ascontiguousarray(np.broadcast_to(np.asarray(values, dtype=np.float32), idx.shape))
        _lib.call("dqnb_logits_set", self._h, idx.ctypes.data, val.ctypes.data, idx.size, self.stream)

    def get(self, indices):
        idx = _i64(np.atleast_1d(indices))
        out = np.empty(idx.shape, dtype=np.float32)
        _lib.call("dqnb_logits_get", self._h, idx.ctypes.data, out.ctypes.data, idx.size, self.stream)
        return out

    def sample(self, uniforms):
        u = np.ascontiguousarray(uniforms, dtype=np.float64)
        out = np.empty(u.shape, dtype=np.int64)
        try:
            _lib.call("dqnb_logits_sample", self._h, u.ctypes.data, u.size, out.ctypes.data, self.stream)
        except _lib.DqnbError as e:
            if "NaN" in str(e):     # what Generator.choice raises for NaN probabilities in the reference
                raise ValueError("probabilities contain NaN") from e
            raise
        return out

    def lse(self):
        out = ctypes.c_float()
        _lib.call("dqnb_logits_lse", self._h, ctypes.byref(out), self.stream)
        return np.float32(out.value)

    def to_host(self):
        out = np.empty(self.capacity, dtype=np.float32)
        _lib.call("dqnb_logits_read", self._h, out.ctypes.data, out.size)
        return out

    def from_host(self, values):
        v = np.ascontiguousarray(values, dtype=np.float32)
        if v.shape != (self.capacity,):
            raise ValueError("logits must have shape (%d,), got %s" % (self.capacity, v.shape))
        _lib.call("dqnb_logits_write", self._h, v.ctypes.data, v.size)


class _LogitsProperty:
    """`_logits` of the reference is a plain ndarray that experiments read and poke; here reading
    copies the device array out and assigning copies a whole array in."""

    def __get__(self, obj, objtype=None):
        return obj._dev.to_host()

    def __set__(self, obj, value):
        obj._dev.from_host(value)


class CircularLogitBuffer:
    """replay_circular.py:148-248."""
    _logits = _LogitsProperty()

    def __init__(self, capacity: int, random_state: np.random.Generator, device=0, stream=0):
        self._dev = _DeviceLogits(capacity, device, stream)
        self._capacity = capacity
        self._size = 0
        self._left_head = 0
        self._right_head = 0
        self._rng_state = random_state

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def size(self) -> int:
        return self._size

    def is_full(self) -> bool:
        return self._size == self._capacity

    def add(self, item: Optional[np.float32] = None) -> None:
        if self.is_full():
            raise BufferError("Buffer is full and cannot be added to. Pop an item first.")
        if item is None:
            self._dev.fill(self._right_head, self._size)   # 0 if empty else lse(all) - log(size)
        else:
            self._dev.set(self._right_head, item)
        self._right_head = self._right_head + 1
        if self._right_head == self._capacity:
            self._right_head = 0
        self._size += 1

    def popleft(self, want_value=True):
        if self._size == 0:
            raise BufferError("Buffer is empty and cannot be popped from. Add an item first.")
        item = self._dev.get(self._left_head)[0] if want_value else None
        self._dev.set(self._left_head, -np.inf)
        self._left_head = self._left_head + 1
        if self._left_head == self._capacity:
            self._left_head = 0
        self._size -= 1
        return item

    def _absolute(self, key):
        key = np.asarray(key)
        if (key >= self._size).any():
            raise KeyError("Buffer is not large enough to index at position %s. Must be in [0,%d)."
                           % (key, self.size - 1))
        return (self._left_head + key) % self._capacity

    def __getitem__(self, key):
        if self._size == 0:
            raise BufferError("Buffer is empty and cannot be indexed. Add an item first.")
        idx = self._absolute(key)
        out = self._dev.get(idx)
        return out.reshape(np.shape(key)) if np.ndim(key) else out[0]

    def __setitem__(self, key, item):
        self._dev.set(self._absolute(key), item)

    def as_probs(self) -> np.ndarray:
        l = self._dev.to_host()
        c = l.max()
        return np.exp(l - (c + np.log(np.sum(np.exp(l - c)))))

    def sample(self, size: int) -> np.ndarray:
        if self._size < size:
            raise BufferError("Cannot sample from buffer with length %d when requested sample size was %d."
                              % (self._size, size))
        absolute_indices = self._dev.sample(self._rng_state.random(size))   # what choice(p=...) draws
        return (absolute_indices - self._left_head) % self._capacity

    def sample_uniform(self, size: int, replace: bool = True) -> np.ndarray:
        if self._size < size:
            raise BufferError("Cannot sample from buffer with length %d when requested sample size was %d."
                              % (self._size, size))
        return self._rng_state.choice(self._size, size=size, replace=replace)

    def get_state(self):
        return {'capacity': self._capacity, 'logits': self._dev.to_host(), 'size': self._size,
                'left_head': self._left_head, 'right_head': self._right_head, 'rng_state': self._rng_state}

    def set_state(self, state) -> None:
        if state['capacity'] != self._capacity:
            self._dev.close()
            self._dev = _DeviceLogits(state['capacity'], self._dev.device, self._dev.stream)
        self._capacity = state['capacity']
        self._dev.from_host(state['logits'])
        self._size = state['size']
        self._left_head = state['left_head']
        self._right_head = state['right_head']
        self._rng_state = state['rng_state']

    def check_valid(self):
        return True, ''


class MGSCReservoirDistribution:
    """replay_circular.py:500-560."""
    _logits = _LogitsProperty()

    def __init__(self, rng_state: np.random.Generator, capacity: int, device=0, stream=0):
        self._capacity = capacity
        self._dev = _DeviceLogits(capacity, device, stream)
        self._size = 0
        self._rng_state = rng_state

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def size(self) -> int:
        return self._size

    def is_full(self) -> bool:
        return self._size == self._capacity

    def add(self, priority: Optional[np.float32] = None) -> None:
        if self.is_full():
            raise BufferError("Buffer is full and cannot be added to. Pop an item first.")
        if priority is None:
            self._dev.fill(self._size, self._size)
        else:
            self._dev.set(self._size, priority)
        self._size += 1

    def replace(self, idx: int, priority: Optional[np.float32] = None) -> None:
        if not self.is_full():
            raise BufferError("Buffer should be full before replacing. Current size=%d while capacity=%d."
                              % (self._size, self._capacity))
        if priority is None:
            self._dev.set(idx, -np.inf)
            self._dev.fill(idx, self._size)
        else:
            self._dev.set(idx, priority)

    def __getitem__(self, key):
        out = self._dev.get(key)
        return out.reshape(np.shape(key)) if np.ndim(key) else out[0]

    def __setitem__(self, key, priority):
        self._dev.set(key, priority)

    def as_probs(self) -> np.ndarray:
        l = self._dev.to_host()
        c = l.max()
        return np.exp(l - (c + np.log(np.sum(np.exp(l - c)))))

    def sample(self, size: int) -> np.ndarray:
        if self._size < size:
            raise BufferError("Cannot sample a batch of size %d from buffer with size %d." % (size, self._size))
        return self._dev.sample(self._rng_state.random(size))

    def sample_uniform(self, size: int, replace: bool = True) -> np.ndarray:
        if self._size < size:
            raise BufferError("Cannot sample a batch of size %d from buffer with size %d." % (size, self._size))
        return self._rng_state.choice(self._size, size=size, replace=replace)

    def get_state(self):
        return {'capacity': self._capacity, 'logits': self._dev.to_host(), 'size': self._size,
                'rng_state': self._rng_state}

    def set_state(self, state) -> None:
        self._capacity = state['capacity']
        self._dev.from_host(state['logits'])
        self._size = state['size']
        self._rng_state = state['rng_state']


class _MgscReplayBase:
    """stack / meta-batch / priority code shared by the two MGSC replays."""

    def get(self, indices):
        for i in indices:
            yield self._storage.get(self._slot(int(i)))

    def slots_of(self, indices) -> np.ndarray:
        """Frame-store slots of replay indices (what the learner gathers from)."""
        return np.asarray([self._slot(int(i)) for i in np.asarray(indices).ravel()], dtype=np.int64)

    def sample_indices(self, size: int) -> np.ndarray:
        return self._distribution.sample(size)

    def sample_slots(self, size: int) -> np.ndarray:
        """Same Generator consumption as sample(); returns store slots for the device path."""
        return self.slots_of(self.sample_indices(size))

    def sample(self, size: int):
        return self.stack_transitions(self.sample_indices(size))

    def stack_transitions(self, indices):
        return self._storage.stack(self.slots_of(indices))

    def batch_of_ids_transitions_and_logits(self, size: int):
        indices = self._distribution.sample_uniform(size, replace=False)
        transitions = self.stack_transitions(indices)
        logits = self._distribution[np.asarray(indices)]
        return indices, transitions, logits

    def batch_of_ids_slots_and_logits(self, size: int):
        """Device-path twin of batch_of_ids_transitions_and_logits: no stacked host copy."""
        indices = self._distribution.sample_uniform(size, replace=False)
        return indices, self.slots_of(indices), self._distribution[np.asarray(indices)]

    def update_priorities(self, indices, priorities):
        self._distribution[np.asarray(indices)] = priorities

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def frame_store(self):
        return self._storage.frames


class MGSCFiFoTransitionReplay(_MgscReplayBase):
    """replay_circular.py:1217-1312: FIFO ring of items + ring of logits."""

    def __init__(self, capacity, structure, random_state: np.random.Generator, encoder=None, decoder=None,
                 device=0, stream=0, initial_planes=0):
        self._capacity = capacity
        self._structure = structure
        self._distribution = CircularLogitBuffer(capacity=capacity, random_state=random_state, device=device,
                                                 stream=stream)
        self._storage = ItemStorage(capacity, structure, encoder, decoder, device, stream, initial_planes)
        self._left_head = self._right_head = self._count = 0
        self._t = 0

    def _slot(self, rel):
        if self._count == 0:
            raise BufferError("Buffer is empty and cannot be indexed. Add an item first.")
        return (self._left_head + rel) % self._capacity

    def add(self, item) -> None:
        if self._count == self._capacity:
            self._distribution.popleft(want_value=False)
            self._left_head = (self._left_head + 1) % self._capacity
            self._count -= 1
        self._distribution.add()
        self._storage.put(self._right_head, item)
        self._right_head = (self._right_head + 1) % self._capacity
        self._count += 1
        self._t += 1

    @property
    def size(self) -> int:
        return self._count

    def transitions_and_logits(self):
        """All stored transitions and the raw logit array (replay_jax.py:940-943)."""
        return self.stack_transitions(np.arange(self.size)), self._distribution._logits

    def get_state(self):
        # CircularBuffer.get_state shape (replay_circular.py:132-140): the full slot list and heads.
        # 'list' is the slot-indexed item list; for frame items a lazy view over one exported snapshot of the HBM store.
        live = np.zeros(self._capacity, dtype=bool)
        live[(self._left_head + np.arange(self._count)) % self._capacity] = True
        snap = self._storage.snapshot()
        if snap is None:
            items = [self._storage.get(slot) if live[slot] else None for slot in range(self._capacity)]
        else:
            items = _uniform.SnapshotSlots(snap, self._capacity, live)
        return {'storage': {'list': items, 'capacity': self._capacity, 'size': self._count,
                            'left_head': self._left_head, 'right_head': self._right_head},
                't': self._t, 'distribution': self._distribution.get_state()}

    def set_state(self, state) -> None:
        st = state['storage']
        if st['capacity'] != self._capacity:
            raise ValueError("snapshot capacity %d != replay capacity %d" % (st['capacity'], self._capacity))
        self._left_head, self._right_head, self._count = st['left_head'], st['right_head'], st['size']
        if isinstance(st['list'], _uniform.SnapshotSlots):
            self._storage.restore(st['list'].snapshot)
        else:
            for k in range(self._count):
                slot = (self._left_head + k) % self._capacity
                self._storage.put(slot, st['list'][slot])
        self._t = state['t']
        self._distribution.set_state(state['distribution'])

    def check_valid(self):
        if self._t < self._count:
            return False, 't should be >= storage size.'
        return self._distribution.check_valid()


class MGSCReservoirTransitionReplay(_MgscReplayBase):
    """replay_circular.py:568-664: Algorithm-R slots + logits (replaced slot re-seeded)."""

    def __init__(self, capacity, structure, random_state: np.random.Generator, encoder=None, decoder=None,
                 device=0, stream=0, initial_planes=0):
        self._capacity = capacity
        self._structure = structure
        self._random_state = random_state
        self._distribution = MGSCReservoirDistribution(random_state, capacity, device=device, stream=stream)
        self._storage = ItemStorage(capacity, structure, encoder, decoder, device, stream, initial_planes)
        self._t = 0

    def _slot(self, idx):
        return idx

    def add(self, item) -> None:
        if self.size == self._capacity:
            j = self._random_state.integers(0, self._t)    # same draw as replay_circular.py:593
            if j < self.size:
                self._storage.put(int(j), item)
                self._distribution.replace(int(j))
        else:
            self._distribution.add()
            self._storage.put(self._t, item)
        self._t += 1

    @property
    def size(self) -> int:
        return self._distribution.size

    def get_state(self):
        snap = self._storage.snapshot()
        if snap is None:
            items = [self._storage.get(k) if k < self.size else None for k in range(self._capacity)]
        else:
            items = _uniform.SnapshotSlots(snap, self._capacity, np.arange(self._capacity) < self.size)
        return {'storage': items,
                't': self._t, 'distribution': self._distribution.get_state(),
                'random_state': self._random_state, 'capacity': self._capacity}

    def set_state(self, state) -> None:
        if isinstance(state['storage'], _uniform.SnapshotSlots):
            self._storage.restore(state['storage'].snapshot)
        else:
            for k, item in enumerate(state['storage']):
                if item is not None:
                    self._storage.put(k, item)
        self._t = state['t']
        self._distribution.set_state(state['distribution'])
        self._random_state = state['random_state']
        self._capacity = state['capacity']

    def check_valid(self):
        return True
