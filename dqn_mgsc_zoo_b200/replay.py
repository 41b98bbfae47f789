"""Drop-in for the uniform replays of dqn_zoo/replay.py, backed by the HBM frame store.

Mirrors (same names, arguments, errors, random-stream consumption):
  Transition                 replay.py:39-44
  UniformDistribution        replay.py:87-160
  TransitionReplay           replay.py:163-243   (FIFO)
  ReservoirTransitionReplay  replay.py:246-333   (Algorithm R)
  TransitionAccumulator      replay.py:1163-1197

Items shaped like `Transition` (two HxWxS uint8 stacks + action, reward, discount) live on
the GPU: frames go to the deduplicated plane pool of `FrameStore`, and `sample()` is one
gather+stack kernel (plus a D2H copy when NumPy arrays are asked for).  The host keeps only
what the reference keeps on the host anyway: the id tables and the NumPy RandomState, which
is shared with the caller and must be consumed draw-for-draw like the reference.  `encoder`/
`decoder` (snappy compression in the reference runners, dqn/run_atari.py:178-190) are
accepted and unused for such items: HBM holds the raw bytes, which is lossless.

Items of any other shape (the reference tests use `Pair(a, b)` of Python ints) are not on
the learner hot path; they are kept in a host list exactly like the reference keeps them.
"""
import collections
import typing
from typing import Any, Optional

import numpy as np

from dqn_mgsc_zoo_b200.frame_store import FrameStore


class Transition(typing.NamedTuple):
    s_tm1: Optional[np.ndarray]
    a_tm1: Optional[int]
    r_t: Optional[float]
    discount_t: Optional[float]
    s_t: Optional[np.ndarray]


def is_frame_item(item) -> bool:
    if len(item) != 5:
        return False
    s0, a, r, d, s1 = item
    ok = (isinstance(s0, np.ndarray) and isinstance(s1, np.ndarray) and s0.dtype == np.uint8
          and s1.dtype == np.uint8 and s0.ndim == 3 and s0.shape == s1.shape and 1 <= s0.shape[2] <= 4)
    return bool(ok and np.ndim(a) == 0 and np.ndim(r) == 0 and np.ndim(d) == 0)


def _scalar_dtype(value, kind):
    """Host dtype of a per-slot scalar column: what np.stack gives the reference for Python scalars
    (int64 actions, float64 rewards / discounts, replay.py:204-205) - never narrower because the FIRST
    item happened to carry an int reward (a synthetic env's `0`), which would truncate later floats."""
    dt = np.asarray(value).dtype
    if kind == 'f':
        return dt if dt.kind == 'f' else np.dtype(np.float64)
    return dt if dt.kind in 'iu' and isinstance(value, np.generic) else np.dtype(np.int64)


class ItemStorage:
    """slot -> item.  Frame items on the device, anything else in a host list."""

    def __init__(self, capacity, structure, encoder=None, decoder=None, device=0, stream=0,
                 initial_planes=0):
        self.capacity, self.structure = capacity, structure
        self.encoder = encoder or (lambda s: s)
        self.decoder = decoder or (lambda s: s)
        self.device, self.stream, self.initial_planes = device, stream, initial_planes
        self.frames = None          # FrameStore once the first frame item arrives
        self.host_items = None      # list once the first generic item arrives
        self.scalars = None         # exact host copies of (a, r, discount) per slot

    def put(self, slot, item):
        if self.frames is None and self.host_items is None:
            if is_frame_item(item):
                h, w, s = item[0].shape
                self.frames = FrameStore(self.capacity, h, w, s, self.device, self.stream,
                                         self.initial_planes)
                self.scalars = [np.zeros(self.capacity, dtype=_scalar_dtype(item[k], kind))
                                for k, kind in ((1, 'i'), (2, 'f'), (3, 'f'))]
            else:
                self.host_items = [None] * self.capacity
        if self.frames is not None:
            s0, a, r, d, s1 = item
            self.frames.put(slot, s0, s1, a, r, d)
            for arr, val in zip(self.scalars, (a, r, d)):
                arr[slot] = val
        else:
            self.host_items[slot] = self.encoder(item)

    def get(self, slot):
        if self.frames is not None:
            s0, _, _, _, s1 = self.frames.gather_host(np.array([slot]))
            a, r, d = (arr[slot].item() for arr in self.scalars)
            return type(self.structure)(s0[0], a, r, d, s1[0])
        return self.decoder(self.host_items[slot])

    def stack(self, slots):
        """The np.stack-per-field batch the reference's sample() returns."""
        slots = np.asarray(slots, dtype=np.int64)
        if self.frames is not None:
            s0, _, _, _, s1 = self.frames.gather_host(slots)
            a, r, d = (arr[slots] for arr in self.scalars)
            return type(self.structure)(s0, a, r, d, s1)
        items = [self.decoder(self.host_items[int(s)]) for s in slots]
        return type(self.structure)(*[np.stack(xs, axis=0) for xs in zip(*items)])


class UniformDistribution:
    """Provides uniform sampling of user-defined integer IDs (replay.py:87-160)."""

    def __init__(self, random_state: np.random.RandomState):
        self._random_state = random_state
        self._ids = []
        self._id_to_index = {}
        # int64 mirror of `_ids` (position -> id), so that sample() is one vectorised take instead of a
        # Python loop over the positions (0.4 ms for the 4 096 draws of a data-parallel step)
        self._ids_np = np.empty(1024, dtype=np.int64)

    def _mirror(self, pos, value):
        if pos >= self._ids_np.shape[0]:
            grown = np.empty(max(2 * self._ids_np.shape[0], pos + 1), dtype=np.int64)
            grown[:self._ids_np.shape[0]] = self._ids_np
            self._ids_np = grown
        self._ids_np[pos] = value

    def add(self, ids) -> None:
        for i in ids:
            if i in self._id_to_index:
                raise IndexError('Cannot add ID %d, it already exists.' % i)
        for i in ids:
            self._id_to_index[i] = len(self._ids)
            self._mirror(len(self._ids), i)
            self._ids.append(i)

    def remove(self, ids) -> None:
        for i in ids:
            if i not in self._id_to_index:
                raise IndexError('Cannot remove ID %d, it does not exist.' % i)
        for i in ids:
            pos = self._id_to_index.pop(i)
            tail = self._ids.pop()
            if tail != i:
                self._ids[pos] = tail
                self._ids_np[pos] = tail
                self._id_to_index[tail] = pos

    def _draw(self, high, size):
        return self._random_state.randint(high, size=size)          # replay.py:121

    def sample(self, size: int) -> np.ndarray:
        positions = self._draw(self.size, size)
        return self._ids_np[positions]

    def ids(self):
        return self._id_to_index.keys()

    @property
    def size(self) -> int:
        return len(self._ids)

    def get_state(self):
        return {'ids': self._ids, 'id_to_index': self._id_to_index}

    def set_state(self, state) -> None:
        self._ids = state['ids']
        self._id_to_index = state['id_to_index']
        self._ids_np = np.empty(max(1024, len(self._ids)), dtype=np.int64)
        self._ids_np[:len(self._ids)] = self._ids

    def check_valid(self):
        if len(self._ids) != len(self._id_to_index):
            return False, 'ids and id_to_index should be the same size.'
        if len(self._ids) != len(set(self._ids)):
            return False, 'IDs should be unique.'
        if len(self._id_to_index.values()) != len(set(self._id_to_index.values())):
            return False, 'Indices should be unique.'
        for i in self._ids:
            if self._ids[self._id_to_index[i]] != i:
                return False, 'ID %d should map to itself.' % i
        return True, ''


class _UniformReplayBase:
    """Shared sampling / state code of the FIFO and reservoir replays."""
    _DISTRIBUTION = UniformDistribution

    def __init__(self, capacity, structure, random_state, encoder=None, decoder=None, device=0,
                 stream=0, initial_planes=0):
        self._capacity = capacity
        self._structure = structure
        self._random_state = random_state
        self._distribution = self._DISTRIBUTION(random_state=random_state)
        self._storage = ItemStorage(capacity, structure, encoder, decoder, device, stream, initial_planes)
        self._live = collections.OrderedDict()   # id -> slot, insertion-ordered like the reference dict
        self._t = 0

    # -- reference API -------------------------------------------------------------
    def get(self, ids):
        for i in ids:
            yield self._storage.get(self._live[i])

    def sample_slots(self, size: int) -> np.ndarray:
        """Draws ids exactly like sample() (same RandomState consumption); returns store slots."""
        ids = self._distribution.sample(size)
        return self._slots_of(ids)

    def sample(self, size: int):
        return self._storage.stack(self.sample_slots(size))

    def ids(self):
        return self._live.keys()

    @property
    def size(self) -> int:
        return len(self._live)

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def frame_store(self) -> Optional[FrameStore]:
        """The HBM store the learner gathers from (None until a frame item was added)."""
        return self._storage.frames

    def get_state(self):
        # Same shape as the reference (replay.py:222-229).  Frame items are read back from
        # HBM (one gather per item): fine for tests/small buffers, ~56 KB per item at scale.
        return {
            'storage': [(i, self._storage.get(slot)) for i, slot in self._live.items()],
            't': self._t,
            'distribution': self._distribution.get_state(),
        }

    def set_state(self, state) -> None:
        self._live = collections.OrderedDict()
        for i, item in state['storage']:
            slot = self._slot_for_id(i)
            self._storage.put(slot, item)
            self._live[i] = slot
        self._t = state['t']
        self._distribution.set_state(state['distribution'])

    def check_valid(self):
        if self._t < len(self._live):
            return False, 't should be >= storage size.'
        if set(self._live.keys()) != set(self._distribution.ids()):
            return False, 'IDs in storage and distribution do not match.'
        return self._distribution.check_valid()

    def _slot_for_id(self, item_id):
        return item_id % self._capacity

    def _slots_of(self, ids):
        return np.asarray(ids, dtype=np.int64) % self._capacity


class TransitionReplay(_UniformReplayBase):
    """Uniform replay with FIFO eviction (replay.py:163-243); ring slot = id mod capacity."""

    def add(self, item) -> None:
        if self.size == self._capacity:
            oldest_id, _ = self._live.popitem(last=False)
            self._distribution.remove([oldest_id])
        item_id = self._t
        self._distribution.add([item_id])
        slot = item_id % self._capacity
        self._storage.put(slot, item)
        self._live[item_id] = slot
        self._t += 1


class ReservoirTransitionReplay(_UniformReplayBase):
    """Uniform replay with Algorithm-R insertion (replay.py:246-333); slot = id."""

    def _draw_replacement(self):
        return self._random_state.randint(0, self._t)   # same draw as replay.py:270

    def add(self, item) -> None:
        if self.size == self._capacity:
            j = self._draw_replacement()
            if j < self.size:
                self._storage.put(int(j), item)
        else:
            item_id = self._t
            self._distribution.add([item_id])
            self._storage.put(item_id, item)
            self._live[item_id] = item_id
        self._t += 1


class TransitionAccumulator:
    """Accumulates timesteps to form transitions (replay.py:1163-1197)."""

    def __init__(self):
        self.reset()

    def step(self, timestep_t, a_t):
        if timestep_t.first():
            self.reset()
        if self._timestep_tm1 is None:
            if not timestep_t.first():
                raise ValueError('Expected FIRST timestep, got %s.' % str(timestep_t))
            self._timestep_tm1, self._a_tm1 = timestep_t, a_t
            return
        transition = Transition(
            s_tm1=self._timestep_tm1.observation, a_tm1=self._a_tm1, r_t=timestep_t.reward,
            discount_t=timestep_t.discount, s_t=timestep_t.observation)
        self._timestep_tm1, self._a_tm1 = timestep_t, a_t
        yield transition

    def reset(self) -> None:
        self._timestep_tm1 = None
        self._a_tm1 = None


def _build_n_step_transition(transitions):
    """Builds a single n-step transition from n 1-step transitions (replay.py:1200-1215)."""
    r_t = 0.0
    discount_t = 1.0
    for transition in transitions:
        r_t += discount_t * transition.r_t
        discount_t *= transition.discount_t
    return Transition(s_tm1=transitions[0].s_tm1, a_tm1=transitions[0].a_tm1, r_t=r_t, discount_t=discount_t,
                      s_t=transitions[-1].s_t)


class NStepTransitionAccumulator:
    """Accumulates timesteps to form n-step transitions (replay.py:1218-1284): nothing on FIRST, one
    n-step transition per MID once n are buffered, all remaining n..1-step ones on LAST."""

    def __init__(self, n):
        self._transitions = collections.deque(maxlen=n)
        self.reset()

    def step(self, timestep_t, a_t):
        if timestep_t.first():
            self.reset()
        if self._timestep_tm1 is None:
            assert self._a_tm1 is None
            if not timestep_t.first():
                raise ValueError('Expected FIRST timestep, got %s.' % str(timestep_t))
            self._timestep_tm1, self._a_tm1 = timestep_t, a_t
            return
        self._transitions.append(Transition(
            s_tm1=self._timestep_tm1.observation, a_tm1=self._a_tm1, r_t=timestep_t.reward,
            discount_t=timestep_t.discount, s_t=timestep_t.observation))
        self._timestep_tm1, self._a_tm1 = timestep_t, a_t
        if timestep_t.last():
            while self._transitions:
                yield _build_n_step_transition(self._transitions)
                self._transitions.popleft()
        else:
            if len(self._transitions) < self._transitions.maxlen:
                return
            yield _build_n_step_transition(self._transitions)

    def reset(self) -> None:
        self._transitions.clear()
        self._timestep_tm1 = None
        self._a_tm1 = None


# The reference runners snappy-compress every stack before it enters the host replay
# (replay.py:1287-1296, dqn/run_atari.py:178-190) to fit 1 M transitions in 25 GB of RAM.  Frame items
# never pass through encoder/decoder here (HBM holds the raw bytes, 7 GB deduplicated), so the codec only
# has to round-trip for callers that still invoke it: snappy when the package exists, else raw bytes.
try:
    import snappy as _snappy
except ImportError:      # not in this image
    _snappy = None


def compress_array(array: np.ndarray):
    """(bytes, shape, dtype) - replay.py:1287-1289."""
    raw = np.ascontiguousarray(array)
    return (_snappy.compress(raw) if _snappy is not None else raw.tobytes()), array.shape, array.dtype


def uncompress_array(compressed) -> np.ndarray:
    """replay.py:1292-1296."""
    compressed_array, shape, dtype = compressed
    byte_string = _snappy.uncompress(compressed_array) if _snappy is not None else compressed_array
    return np.frombuffer(byte_string, dtype=dtype).reshape(shape)
