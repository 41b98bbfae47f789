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
import collections.abc
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


class FrameSnapshot:
    """Host copy of a whole HBM frame store (one `dqnb_replay_export` blob) + the exact host scalars, from which
    single items are materialised on demand.  Picklable as is (NumPy arrays): a 1 M-transition replay is ~8 GB
    of deduplicated planes instead of 1 M gathered 56 KB items."""

    FIELDS = ("magic", "version", "capacity", "H", "W", "S", "stride", "pool", "nfree", "have_last", "off_slot_ids",
              "off_actions", "off_rewards", "off_discounts", "off_refcount", "off_free", "off_last_ids",
              "off_last_st", "off_planes", "total")

    def __init__(self, blob, scalars, structure_type):
        self.blob, self.scalars, self.structure_type = blob, scalars, structure_type

    def _header(self):
        h = np.frombuffer(self.blob, dtype=np.int64, count=len(self.FIELDS))
        return dict(zip(self.FIELDS, (int(x) for x in h)))

    def item(self, slot):
        """The item stored in `slot`, rebuilt from the snapshot's planes (plane id -1 = the all-zero plane)."""
        f = self._header()
        ids = np.frombuffer(self.blob, dtype=np.int32, count=8, offset=f["off_slot_ids"] + 32 * int(slot))
        H, W, S = f["H"], f["W"], f["S"]

        def stack(plane_ids):
            out = np.zeros((H, W, S), dtype=np.uint8)
            for c, p in enumerate(plane_ids):
                if p >= 0:
                    out[:, :, c] = np.frombuffer(self.blob, dtype=np.uint8, count=H * W,
                                                 offset=f["off_planes"] + int(p) * f["stride"]).reshape(H, W)
            return out
        a, r, d = (arr[int(slot)].item() for arr in self.scalars)
        return self.structure_type(stack(ids[:S]), a, r, d, stack(ids[S:2 * S]))


class SnapshotPairs(collections.abc.Sequence):
    """`[(id, item), ...]` of the reference's get_state()['storage'] (replay.py:224), materialised lazily."""

    def __init__(self, snapshot, ids, slots):
        self.snapshot, self.ids, self.slots = snapshot, list(ids), list(slots)

    def __len__(self):
        return len(self.ids)

    def __getitem__(self, k):
        if isinstance(k, slice):
            return [self[i] for i in range(*k.indices(len(self)))]
        return self.ids[k], self.snapshot.item(self.slots[k])


class SnapshotSlots(collections.abc.Sequence):
    """Slot-indexed item list (`CircularBuffer._list`, replay_circular.py:132-140; reservoir storage list):
    entry = item or None for slots that hold nothing."""

    def __init__(self, snapshot, capacity, live):
        self.snapshot, self.capacity, self.live = snapshot, capacity, live     # live: bool array per slot

    def __len__(self):
        return self.capacity

    def __getitem__(self, k):
        if isinstance(k, slice):
            return [self[i] for i in range(*k.indices(len(self)))]
        if k < 0:
            k += self.capacity
        return self.snapshot.item(k) if self.live[k] else None


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
            sc = self.scalars
            sc[0][slot] = a
            sc[1][slot] = r
            sc[2][slot] = d
        else:
            self.host_items[slot] = self.encoder(item)

    def get(self, slot):
        if self.frames is not None:
            s0, _, _, _, s1 = self.frames.gather_host(np.array([slot]))
            a, r, d = (arr[slot].item() for arr in self.scalars)
            return type(self.structure)(s0[0], a, r, d, s1[0])
        return self.decoder(self.host_items[slot])

    def snapshot(self):
        """FrameSnapshot of the device store (None when the items live on the host)."""
        if self.frames is None:
            return None
        return FrameSnapshot(self.frames.export_blob(), [a.copy() for a in self.scalars], type(self.structure))

    def restore(self, snapshot):
        """Inverse of snapshot(): one H2D pass instead of one put per item."""
        f = snapshot._header()
        if self.frames is None:
            self.frames = FrameStore(self.capacity, f["H"], f["W"], f["S"], self.device, self.stream,
                                     self.initial_planes)
            self.host_items = None
        self.frames.import_blob(snapshot.blob)
        self.scalars = [a.copy() for a in snapshot.scalars]

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
        """Same shape as the reference (replay.py:222-229): 'storage' is a sequence of (id, item).  For frame
        items it is a lazy view over ONE exported snapshot of the HBM store (dqnb_replay_export: tables + plane
        pool at PCIe rate); items are materialised only if somebody indexes it.  Pickles compactly."""
        snap = self._storage.snapshot()
        if snap is None:
            storage = [(i, self._storage.get(slot)) for i, slot in self._live.items()]
        else:
            storage = SnapshotPairs(snap, self._live.keys(), self._live.values())
        return {'storage': storage, 't': self._t, 'distribution': self._distribution.get_state()}

    def set_state(self, state) -> None:
        storage = state['storage']
        self._live = collections.OrderedDict()
        if isinstance(storage, SnapshotPairs):          # our own checkpoint: one import, no per-item puts
            self._storage.restore(storage.snapshot)
            self._live.update(zip(storage.ids, storage.slots))
        else:                                           # a reference-shaped list of (id, item)
            for i, item in storage:
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
        live, dist = self._live, self._distribution
        item_id = self._t
        if type(dist) is UniformDistribution:
            # the stock distribution: remove([oldest]) + add([new]) of replay.py:87-160 written out for ONE id (same
            # state afterwards; the list wrappers, loops and property calls were a third of an add's host time)
            ids, index = dist._ids, dist._id_to_index
            if len(live) == self._capacity:
                oldest_id, _ = live.popitem(last=False)
                if oldest_id not in index:
                    raise IndexError('Cannot remove ID %d, it does not exist.' % oldest_id)
                pos = index.pop(oldest_id)
                tail = ids.pop()
                if tail != oldest_id:
                    ids[pos] = tail
                    dist._ids_np[pos] = tail
                    index[tail] = pos
            if item_id in index:
                raise IndexError('Cannot add ID %d, it already exists.' % item_id)
            n = len(ids)
            index[item_id] = n
            if n < dist._ids_np.shape[0]:
                dist._ids_np[n] = item_id
            else:
                dist._mirror(n, item_id)
            ids.append(item_id)
        else:
            if len(live) == self._capacity:
                oldest_id, _ = live.popitem(last=False)
                dist.remove([oldest_id])
            dist.add([item_id])
        slot = item_id % self._capacity
        self._storage.put(slot, item)
        live[item_id] = slot
        self._t = item_id + 1


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


# ------------------------------------------------------------------------------------------------
# Prioritized replay: replay.py:336-784 (importance weights, SumTree,
# PrioritizedDistribution) and :1046-1160 (PrioritizedTransitionReplay).
# ------------------------------------------------------------------------------------------------
def _power(base, exponent) -> np.ndarray:
    """Same as usual power except `0 ** 0` is zero (replay.py:336-341)."""
    base = np.asarray(base)
    return np.where(base == 0.0, 0.0, base ** exponent)


def importance_sampling_weights(probabilities: np.ndarray, uniform_probability: float, exponent: float,
                                normalize: bool) -> np.ndarray:
    """Importance sampling weights from sampling probabilities (replay.py:344-376)."""
    if not 0.0 <= exponent <= 1.0:
        raise ValueError('Require 0 <= exponent <= 1.')
    if not 0.0 <= uniform_probability <= 1.0:
        raise ValueError('Expected 0 <= uniform_probability <= 1.')
    weights = (uniform_probability / probabilities) ** exponent
    if normalize:
        weights /= np.max(weights)
    if not np.isfinite(weights).all():
        raise ValueError('Weights are not finite: %s.' % weights)
    return weights


class SumTree:
    """replay.py:379-559, with the tree itself in device memory (`dqnb_sumtree_*`, csrc/sumtree.cu).

    The device array has the reference's layout (2 * capacity float64, node i = left + right, root at 1) and is
    updated / queried with the reference's float64 arithmetic, so nodes and query results are bit-identical; what
    the reference does in Python loops - 20 levels per index of a `set`, 20 per target of a `query` at capacity
    1 M - is one small kernel each.  A host mirror of the LEAVES serves `get` / `values` without a device round
    trip (the caller hands every value through the host anyway)."""

    def __init__(self, device=0, stream=0):
        self._size = 0
        self._first_leaf = 0
        self._device, self._stream = device, stream
        self._h = None
        self._leaves = np.zeros(0, dtype=np.float64)     # mirror of storage[first_leaf:]
        self._nonzero = 0                                # leaves != 0: root == 0.0 iff this is 0 (values are >= 0)

    def __del__(self):
        try:
            self._free()
        except Exception:
            pass

    def _free(self):
        if self._h is not None:
            from dqn_mgsc_zoo_b200 import _lib
            _lib.call("dqnb_sumtree_destroy", self._h)
            self._h = None

    def resize(self, size: int) -> None:
        self._initialize(size, values=None)

    def get(self, indices) -> np.ndarray:
        indices = np.asarray(indices)
        if not ((0 <= indices) & (indices < self.size)).all():
            raise IndexError('index out of range, expect 0 <= index < %s' % self.size)
        return self.values[indices]

    def set(self, indices, values) -> None:
        values = np.asarray(values)
        if not np.isfinite(values).all() or (values < 0.0).any():
            raise ValueError('value must be finite and positive.')
        idx = np.ascontiguousarray(np.asarray(indices).ravel(), dtype=np.int64)
        val = np.ascontiguousarray(np.broadcast_to(np.asarray(values, dtype=np.float64).ravel(), idx.shape))
        if idx.size == 0:
            return
        if not ((0 <= idx) & (idx < self._first_leaf)).all():
            raise IndexError('index out of range, expect 0 <= index < %s' % self._first_leaf)
        touched = np.unique(idx) if idx.size > 1 else idx
        before = np.count_nonzero(self._leaves[touched])
        self._leaves[idx] = val                      # duplicates: the last one wins, as in `self.values[indices] = values`
        self._nonzero += np.count_nonzero(self._leaves[touched]) - before
        from dqn_mgsc_zoo_b200 import _lib
        _lib.call("dqnb_sumtree_set", self._h, idx.ctypes.data, val.ctypes.data, idx.size, self._stream)

    def set_all(self, values) -> None:
        values = np.asarray(values)
        if not np.isfinite(values).all() or (values < 0.0).any():
            raise ValueError('Values must be finite positive numbers.')
        self._initialize(len(values), values)

    def _device_query(self, x, scale_by_root):
        import ctypes
        from dqn_mgsc_zoo_b200 import _lib
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty(x.shape, dtype=np.int64)
        root = ctypes.c_double()
        try:
            _lib.call("dqnb_sumtree_query", self._h, x.ctypes.data, x.size, int(scale_by_root), out.ctypes.data,
                      ctypes.byref(root), self._stream)
        except _lib.DqnbError as e:
            if "Require 0 <= target" in str(e):
                raise ValueError('Require 0 <= target < total sum.') from e
            raise
        return out, root.value

    def query(self, targets):
        """Smallest indices where `target <` cumulative value sum up to index (replay.py:424-440)."""
        if self._h is None:
            raise ValueError('Require 0 <= target < total sum.')
        return [int(i) for i in self._device_query(targets, 0)[0]]

    def query_uniform(self, uniforms):
        """-> (indices for targets uniform * root, root): the sampling step of PrioritizedDistribution.sample
        (replay.py:697-699) with the multiplication done on the device, same float64 product."""
        return self._device_query(uniforms, 1)

    def root(self) -> float:
        if self.size == 0 or self._h is None:
            return np.nan
        return self._device_query(np.zeros(0), 0)[1]

    @property
    def values(self) -> np.ndarray:
        return self._leaves[:self._size]

    @property
    def size(self) -> int:
        return self._size

    @property
    def capacity(self) -> int:
        return self._first_leaf

    def _read_storage(self):
        from dqn_mgsc_zoo_b200 import _lib
        storage = np.zeros(2 * self._first_leaf, dtype=np.float64)
        if self._h is not None:
            _lib.call("dqnb_sumtree_read", self._h, storage.ctypes.data, storage.size)
        return storage

    def get_state(self):
        return {'size': self._size, 'storage': self._read_storage(), 'first_leaf': self._first_leaf}

    def set_state(self, state) -> None:
        self._size = state['size']
        self._upload(np.array(state['storage'], dtype=np.float64), state['first_leaf'])

    def check_valid(self):
        storage = self._read_storage()
        if len(storage) != 2 * self._first_leaf:
            return False, 'first_leaf should be half the size of storage.'
        if not 0 <= self.size <= self.capacity:
            return False, 'Require 0 <= self.size <= self.capacity.'
        if len(self.values) != self.size:
            return False, 'Number of values should be equal to the size.'
        n = self._first_leaf
        if n > 1:
            sums = storage[2:2 * n:2] + storage[3:2 * n:2]
            bad = np.nonzero(storage[1:n] != sums)[0]
            if bad.size:
                return False, 'Non-leaf node %d should be sum of child nodes.' % (bad[0] + 1)
        if not np.array_equal(storage[n:], self._leaves):
            return False, 'Host mirror of the leaves differs from the device tree.'
        return True, ''

    def _upload(self, storage, first_leaf):
        """Makes the device tree equal to `storage` (2 * first_leaf doubles); (re)allocates on a capacity change."""
        import ctypes
        from dqn_mgsc_zoo_b200 import _lib
        if first_leaf != self._first_leaf or self._h is None:
            self._free()
            self._first_leaf = first_leaf
            if first_leaf > 0:
                h = ctypes.c_void_p()
                _lib.call("dqnb_sumtree_create", int(first_leaf), self._device, ctypes.byref(h))
                self._h = h
        if first_leaf > 0:
            storage = np.ascontiguousarray(storage, dtype=np.float64)
            _lib.call("dqnb_sumtree_write", self._h, storage.ctypes.data, storage.size)
        self._leaves = np.array(storage[first_leaf:], dtype=np.float64)
        self._nonzero = int(np.count_nonzero(self._leaves))

    def _initialize(self, size: int, values) -> None:
        """Resizes storage and sets new values if supplied (replay.py:495-528)."""
        assert size >= 0
        assert values is None or len(values) == size
        if size < self.size:
            new_values = self.values[:size] if values is None else values
            self._size = size
            self._set_values(new_values, self._first_leaf)
        elif size <= self.capacity:
            self._size = size
            if values is not None:
                self._set_values(values, self._first_leaf)
        else:
            new_capacity = 1
            while new_capacity < size:
                new_capacity *= 2
            new_values = np.array(self.values) if values is None else values
            self._size = size
            self._set_values(new_values, new_capacity)

    def _set_values(self, values, first_leaf) -> None:
        """replay.py:530-540: leaves, zero tail, every sum node - level by level, the same additions."""
        values = np.asarray(values, dtype=np.float64)
        assert len(values) <= first_leaf
        storage = np.zeros(2 * first_leaf, dtype=np.float64)
        storage[first_leaf:first_leaf + len(values)] = values
        level = first_leaf // 2
        while level >= 1:
            storage[level:2 * level] = storage[2 * level:4 * level:2] + storage[2 * level + 1:4 * level:2]
            level //= 2
        self._upload(storage, first_leaf)


class PrioritizedDistribution:
    """Distribution for weighted sampling of user-defined integer IDs (replay.py:562-784)."""

    def __init__(self, priority_exponent: float, uniform_sample_probability: float, random_state,
                 min_capacity: int = 0, max_capacity: Optional[int] = None, device=0, stream=0):
        if priority_exponent < 0.0:
            raise ValueError('Require priority_exponent >= 0.')
        self._priority_exponent = priority_exponent
        if not 0.0 <= uniform_sample_probability <= 1.0:
            raise ValueError('Require 0 <= uniform_sample_probability <= 1.')
        if max_capacity is not None and max_capacity < min_capacity:
            raise ValueError('Require max_capacity >= min_capacity.')
        if min_capacity < 0:
            raise ValueError('Require min_capacity >= 0.')
        self._uniform_sample_probability = uniform_sample_probability
        self._max_capacity = max_capacity
        self._sum_tree = SumTree(device, stream)
        self._sum_tree.resize(min_capacity)
        self._random_state = random_state
        self._id_to_index = {}
        self._index_to_id = {}
        self._inactive_indices = list(range(min_capacity))
        self._active_indices = []
        self._active_indices_location = {}
        # int64 mirrors for vectorised lookups in sample(): position -> active index, index -> id
        self._active_np = np.empty(max(16, min_capacity), dtype=np.int64)
        self._index_to_id_np = np.full(max(16, min_capacity), -1, dtype=np.int64)

    def _draw_positions(self, size):
        return self._random_state.randint(self.size, size=size)          # replay.py:686

    @staticmethod
    def _grown(arr, n, fill):
        if n <= arr.shape[0]:
            return arr
        out = np.full(max(n, 2 * arr.shape[0]), fill, dtype=arr.dtype)
        out[:arr.shape[0]] = arr
        return out

    def ensure_capacity(self, capacity: int) -> None:
        if self._max_capacity is not None and capacity > self._max_capacity:
            raise ValueError('capacity %d cannot exceed max_capacity %d' % (capacity, self._max_capacity))
        if capacity <= self._sum_tree.size:
            return
        self._inactive_indices.extend(range(self._sum_tree.size, capacity))
        self._sum_tree.resize(capacity)

    def add_priorities(self, ids, priorities) -> None:
        for i in ids:
            if i in self._id_to_index:
                raise IndexError('ID %d already exists.' % i)
        new_size = self.size + len(ids)
        if self._max_capacity is not None and new_size > self._max_capacity:
            raise ValueError('Cannot add IDs as max capacity would be exceeded.')
        if new_size > self.capacity:
            candidate_capacity = max(new_size, 2 * self.capacity)
            new_capacity = candidate_capacity if self._max_capacity is None else min(self._max_capacity,
                                                                                    candidate_capacity)
            self.ensure_capacity(new_capacity)
        self._active_np = self._grown(self._active_np, new_size, 0)
        self._index_to_id_np = self._grown(self._index_to_id_np, self._sum_tree.size, -1)
        indices = []
        for i in ids:
            idx = self._inactive_indices.pop()
            self._active_indices_location[idx] = len(self._active_indices)
            self._active_np[len(self._active_indices)] = idx
            self._active_indices.append(idx)
            self._id_to_index[i] = idx
            self._index_to_id[idx] = i
            self._index_to_id_np[idx] = i
            indices.append(idx)
        self._sum_tree.set(indices, _power(priorities, self._priority_exponent))

    def remove_priorities(self, ids) -> None:
        indices = []
        for i in ids:
            try:
                idx = self._id_to_index[i]
            except IndexError as err:
                raise IndexError('Cannot remove ID %d, it does not exist.' % i) from err
            indices.append(idx)
        for i, idx in zip(ids, indices):
            del self._id_to_index[i]
            del self._index_to_id[idx]
            self._index_to_id_np[idx] = -1
            j = self._active_indices_location[idx]
            self._active_indices[j], self._active_indices[-1] = self._active_indices[-1], self._active_indices[j]
            self._active_np[j] = self._active_indices[j]
            self._active_indices_location[self._active_indices[j]] = j
            self._active_indices_location.pop(self._active_indices.pop())
        self._inactive_indices.extend(indices)
        self._sum_tree.set(indices, np.zeros((len(indices),), dtype=np.float64))

    def update_priorities(self, ids, priorities) -> None:
        indices = []
        for i in ids:
            if i not in self._id_to_index:
                raise IndexError('ID %d does not exist.' % i)
            indices.append(self._id_to_index[i])
        self._sum_tree.set(indices, _power(priorities, self._priority_exponent))

    def sample_indices(self, size: int):
        """The draws of sample() (replay.py:681-717) -> (sum-tree indices, sample probabilities); consumes the
        shared RandomState exactly like the reference: randint, then uniform only if the tree sum is non-zero,
        then uniform for the uniform/prioritized mix."""
        if self.size == 0:
            raise RuntimeError('No IDs to sample.')
        uniform_indices = self._active_np[self._draw_positions(size)]
        root = 0.0
        if self._sum_tree._nonzero == 0:          # root() == 0.0: every priority is zero
            prioritized_indices = uniform_indices
        else:
            prioritized_indices, root = self._sum_tree.query_uniform(self._random_state.uniform(size=size))
        usp = self._uniform_sample_probability
        indices = np.where(self._random_state.uniform(size=size) < usp, uniform_indices, prioritized_indices)
        uniform_prob = np.asarray(1.0 / self.size)
        priorities = self._sum_tree.get(indices)
        if root == 0.0:
            prioritized_probs = np.full_like(priorities, fill_value=uniform_prob)
        else:
            prioritized_probs = priorities / root
        sample_probs = (1.0 - usp) * prioritized_probs + usp * uniform_prob
        return indices, sample_probs

    def sample(self, size: int):
        """Returns sample of IDs with corresponding probabilities."""
        indices, sample_probs = self.sample_indices(size)
        return self._index_to_id_np[indices], sample_probs

    def get_exponentiated_priorities(self, ids):
        indices = np.fromiter((self._id_to_index[i] for i in ids), dtype=np.int64, count=len(ids))
        return self._sum_tree.get(indices)

    def ids(self):
        return self._id_to_index.keys()

    @property
    def capacity(self) -> int:
        return self._sum_tree.size

    @property
    def size(self) -> int:
        return len(self._id_to_index)

    def get_state(self):
        return {'sum_tree': self._sum_tree.get_state(), 'id_to_index': self._id_to_index,
                'index_to_id': self._index_to_id, 'inactive_indices': self._inactive_indices,
                'active_indices': self._active_indices, 'active_indices_location': self._active_indices_location}

    def set_state(self, state) -> None:
        self._sum_tree.set_state(state['sum_tree'])
        self._id_to_index = state['id_to_index']
        self._index_to_id = state['index_to_id']
        self._inactive_indices = state['inactive_indices']
        self._active_indices = state['active_indices']
        self._active_indices_location = state['active_indices_location']
        self._active_np = np.empty(max(16, self._sum_tree.size), dtype=np.int64)
        self._active_np[:len(self._active_indices)] = self._active_indices
        self._index_to_id_np = np.full(max(16, self._sum_tree.size), -1, dtype=np.int64)
        for idx, i in self._index_to_id.items():
            self._index_to_id_np[idx] = i

    def check_valid(self):
        if len(self._id_to_index) != len(self._index_to_id):
            return False, 'ID to index maps are not the same size.'
        for i in self._id_to_index:
            if self._index_to_id[self._id_to_index[i]] != i:
                return False, 'ID %d should map to itself.' % i
        if len(set(self._inactive_indices)) != len(self._inactive_indices):
            return False, 'Inactive indices should be unique.'
        if len(set(self._active_indices)) != len(self._active_indices):
            return False, 'Active indices should be unique.'
        if set(self._active_indices) != set(self._index_to_id.keys()):
            return False, 'Active indices should match index to ID mapping keys.'
        all_indices = self._inactive_indices + self._active_indices
        if sorted(all_indices) != list(range(self._sum_tree.size)):
            return False, 'Inactive and active indices should partition all indices.'
        if len(self._active_indices) != len(self._active_indices_location):
            return False, 'Active indices and their location should be the same size.'
        for j, i in enumerate(self._active_indices):
            if j != self._active_indices_location[i]:
                return False, 'Active index location %d not correct for index %d.' % (j, i)
        return self._sum_tree.check_valid()


class PrioritizedTransitionReplay:
    """Prioritized replay, proportional variant (replay.py:1046-1160): FIFO storage in the HBM frame store
    (ring slot = id mod capacity, as TransitionReplay), priorities in the device sum tree."""
    _DISTRIBUTION = PrioritizedDistribution

    def __init__(self, capacity: int, structure, priority_exponent: float, importance_sampling_exponent,
                 uniform_sample_probability: float, normalize_weights: bool, random_state, encoder=None,
                 decoder=None, device=0, stream=0, initial_planes=0):
        self._capacity = capacity
        self._structure = structure
        self._random_state = random_state
        self._distribution = self._DISTRIBUTION(
            min_capacity=capacity, max_capacity=capacity, priority_exponent=priority_exponent,
            uniform_sample_probability=uniform_sample_probability, random_state=random_state, device=device,
            stream=stream)
        self._importance_sampling_exponent = importance_sampling_exponent
        self._normalize_weights = normalize_weights
        self._storage = ItemStorage(capacity, structure, encoder, decoder, device, stream, initial_planes)
        self._live = collections.OrderedDict()     # id -> slot
        self._t = 0

    def add(self, item, priority: float) -> None:
        if self.size == self._capacity:
            oldest_id, _ = self._live.popitem(last=False)
            self._distribution.remove_priorities([oldest_id])
        item_id = self._t
        self._distribution.add_priorities([item_id], [priority])
        slot = item_id % self._capacity
        self._storage.put(slot, item)
        self._live[item_id] = slot
        self._t += 1

    def get(self, ids):
        for i in ids:
            yield self._storage.get(self._live[i])

    def sample_slots_ids_weights(self, size: int):
        """The draws and weights of sample() without the gather: (store slots, ids, importance weights)."""
        ids, probabilities = self._distribution.sample(size)
        weights = importance_sampling_weights(probabilities, uniform_probability=1.0 / self.size,
                                              exponent=self.importance_sampling_exponent,
                                              normalize=self._normalize_weights)
        return np.asarray(ids, dtype=np.int64) % self._capacity, ids, weights

    def sample(self, size: int):
        slots, ids, weights = self.sample_slots_ids_weights(size)
        return self._storage.stack(slots), ids, weights

    def update_priorities(self, ids, priorities) -> None:
        priorities = np.asarray(priorities)
        self._distribution.update_priorities(ids, priorities)

    @property
    def size(self) -> int:
        return len(self._live)

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def importance_sampling_exponent(self):
        return self._importance_sampling_exponent(self._t)

    @property
    def frame_store(self) -> Optional[FrameStore]:
        return self._storage.frames

    get_state = _UniformReplayBase.get_state
    set_state = _UniformReplayBase.set_state

    def _slot_for_id(self, item_id):
        return item_id % self._capacity

    def check_valid(self):
        if self._t < len(self._live):
            return False, 't should be >= storage size.'
        if set(self._live.keys()) != set(self._distribution.ids()):
            return False, 'IDs in storage and distribution do not match.'
        return self._distribution.check_valid()
