"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference replay semantics.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
arm may import this module.  It restates, in compact array form, what the
reference's host-side replay classes compute so that the CUDA replay store can
be checked for bit-exact indices, batches and reservoir decisions:

  UniformIds          <- replay.py:87-160   (UniformDistribution)
  FifoReplay          <- replay.py:163-243  (TransitionReplay)
  ReservoirReplay     <- replay.py:246-333  (ReservoirTransitionReplay)
  lse / softmax_probs <- replay_circular.py:69-77
  RingLogits          <- replay_circular.py:148-248 (CircularLogitBuffer)
  MgscFifoReplay      <- replay_circular.py:1217-1312
  MgscReservoirReplay <- replay_circular.py:500-664
  PairAccumulator     <- replay.py:1163-1197 (TransitionAccumulator)

Parity is pinned: tests/test_oracle_replay.py runs each class against the
unmodified reference module (oracle/ref_shim.py) on identical seeds/call
sequences, and against the committed fixtures in tests/golden/.
"""
import collections

import numpy as np

Transition = collections.namedtuple("Transition", "s_tm1 a_tm1 r_t discount_t s_t")


def _stack_fields(structure, items):
    cols = zip(*items)
    return type(structure)(*[np.stack(c, axis=0) for c in cols])


class UniformIds:
    """Position->id table with swap-with-last removal (replay.py:95-125)."""

    def __init__(self, random_state):
        self.rs = random_state
        self.table = []          # position -> id
        self.where = {}          # id -> position

    def add(self, ids):
        for i in ids:
            if i in self.where:
                raise IndexError("Cannot add ID %d, it already exists." % i)
        for i in ids:
            self.where[i] = len(self.table)
            self.table.append(i)

    def remove(self, ids):
        for i in ids:
            if i not in self.where:
                raise IndexError("Cannot remove ID %d, it does not exist." % i)
        for i in ids:
            pos = self.where.pop(i)
            tail = self.table.pop()
            if tail != i:           # the last id takes the freed position
                self.table[pos] = tail
                self.where[tail] = pos

    @property
    def size(self):
        return len(self.table)

    def sample(self, n):
        # replay.py:121 draws with RandomState.randint, the replay_circular.py:285 twin with Generator.integers
        pos = self.rs.integers(self.size, size=n) if isinstance(self.rs, np.random.Generator) \
            else self.rs.randint(self.size, size=n)
        return np.array([self.table[p] for p in pos], dtype=np.int64)


class FifoReplay:
    """Evict-oldest store keyed by a running id (replay.py:184-206)."""

    def __init__(self, capacity, structure, random_state):
        self.capacity, self.structure = capacity, structure
        self.dist = UniformIds(random_state)
        self.items = collections.OrderedDict()
        self.t = 0

    @property
    def size(self):
        return len(self.items)

    def add(self, item):
        if len(self.items) == self.capacity:
            old, _ = self.items.popitem(last=False)
            self.dist.remove([old])
        self.dist.add([self.t])
        self.items[self.t] = item
        self.t += 1

    def sample_ids(self, n):
        return self.dist.sample(n)

    def gather(self, ids):
        return _stack_fields(self.structure, [self.items[i] for i in ids])

    def sample(self, n):
        return self.gather(self.sample_ids(n))


class ReservoirReplay:
    """Algorithm R with j ~ randint(0, t) once full (replay.py:267-283)."""

    def __init__(self, capacity, structure, random_state):
        self.capacity, self.structure, self.rs = capacity, structure, random_state
        self.dist = UniformIds(random_state)
        self.items = {}
        self.t = 0
        self.last_decision = None   # slot replaced, -1 dropped, or fill slot

    @property
    def size(self):
        return len(self.items)

    def add(self, item):
        if len(self.items) == self.capacity:
            j = int(self.rs.integers(0, self.t) if isinstance(self.rs, np.random.Generator)
                    else self.rs.randint(0, self.t))       # replay.py:270 / replay_circular.py:436
            if j < self.capacity:
                self.items[j] = item
                self.last_decision = j
            else:
                self.last_decision = -1
        else:
            self.dist.add([self.t])
            self.items[self.t] = item
            self.last_decision = self.t
        self.t += 1

    def sample_ids(self, n):
        return self.dist.sample(n)

    def gather(self, ids):
        return _stack_fields(self.structure, [self.items[i] for i in ids])

    def sample(self, n):
        return self.gather(self.sample_ids(n))


def lse(x):
    """max-shifted logsumexp in the array's own dtype (replay_circular.py:74-77)."""
    c = x.max()
    return c + np.log(np.sum(np.exp(x - c)))


def softmax_probs(x):
    return np.exp(x - lse(x))


class RingLogits:
    """float32 logit ring, -inf marks empty slots (replay_circular.py:148-228)."""

    def __init__(self, capacity, rng):
        self.l = np.full((capacity,), -np.inf, dtype=np.float32)
        self.capacity, self.size, self.head, self.tail, self.rng = capacity, 0, 0, 0, rng

    def push(self, value=None):
        if self.size == self.capacity:
            raise BufferError("Buffer is full and cannot be added to. Pop an item first.")
        if value is None:
            value = 0 if self.size == 0 else lse(self.l) - np.log(self.size)
        self.l[self.tail] = value
        self.tail = (self.tail + 1) % self.capacity
        self.size += 1

    def popleft(self):
        if self.size == 0:
            raise BufferError("Buffer is empty and cannot be popped from. Add an item first.")
        v = self.l[self.head]
        self.l[self.head] = -np.inf
        self.head = (self.head + 1) % self.capacity
        self.size -= 1
        return v

    def _abs(self, rel):
        rel = np.asarray(rel)
        if (rel >= self.size).any():
            raise KeyError("Buffer is not large enough to index at position %s." % (rel,))
        return (self.head + rel) % self.capacity

    def get(self, rel):
        if self.size == 0:
            raise BufferError("Buffer is empty and cannot be indexed. Add an item first.")
        return self.l[self._abs(rel)]

    def set(self, rel, values):
        self.l[self._abs(rel)] = values

    def sample(self, n):
        if self.size < n:
            raise BufferError("Cannot sample %d from buffer with length %d." % (n, self.size))
        absolute = self.rng.choice(self.capacity, size=n, p=softmax_probs(self.l))
        return (absolute - self.head) % self.capacity

    def sample_uniform(self, n, replace=True):
        if self.size < n:
            raise BufferError("Cannot sample %d from buffer with length %d." % (n, self.size))
        return self.rng.choice(self.size, size=n, replace=replace)


class MgscFifoReplay:
    """FIFO ring + logit ring (replay_circular.py:1239-1279)."""

    def __init__(self, capacity, structure, rng):
        self.capacity, self.structure = capacity, structure
        self.logits = RingLogits(capacity, rng)
        self.ring = [None] * capacity
        self.head = self.tail = self.count = 0
        self.t = 0

    @property
    def size(self):
        return self.count

    def add(self, item):
        if self.count == self.capacity:
            self.logits.popleft()
            self.head = (self.head + 1) % self.capacity
            self.count -= 1
        self.logits.push()
        self.ring[self.tail] = item
        self.tail = (self.tail + 1) % self.capacity
        self.count += 1
        self.t += 1

    def stack_transitions(self, rel):
        return _stack_fields(self.structure,
                             [self.ring[(self.head + int(k)) % self.capacity] for k in rel])

    def sample_indices(self, n):
        return self.logits.sample(n)

    def sample(self, n):
        return self.stack_transitions(self.sample_indices(n))

    def batch_of_ids_transitions_and_logits(self, n):
        rel = self.logits.sample_uniform(n, replace=False)
        return rel, self.stack_transitions(rel), self.logits.get(np.asarray(rel))

    def update_priorities(self, rel, values):
        self.logits.set(np.asarray(rel), values)


class MgscReservoirReplay:
    """Reservoir slots + logits; replaced slot re-seeded (replay_circular.py:518-533,591-606)."""

    def __init__(self, capacity, structure, rng):
        self.capacity, self.structure, self.rng = capacity, structure, rng
        self.l = np.full((capacity,), -np.inf, dtype=np.float32)
        self.slots = [None] * capacity
        self.count = 0
        self.t = 0
        self.last_decision = None

    @property
    def size(self):
        return self.count

    def add(self, item):
        if self.count == self.capacity:
            j = int(self.rng.integers(0, self.t))
            if j < self.capacity:
                self.slots[j] = item
                self.l[j] = -np.inf
                self.l[j] = lse(self.l) - np.log(self.count)
                self.last_decision = j
            else:
                self.last_decision = -1
        else:
            self.l[self.count] = 0 if self.count == 0 else lse(self.l) - np.log(self.count)
            self.count += 1
            self.slots[self.t] = item
            self.last_decision = self.t
        self.t += 1

    def stack_transitions(self, idx):
        return _stack_fields(self.structure, [self.slots[int(k)] for k in idx])

    def sample_indices(self, n):
        if self.count < n:
            raise BufferError("Cannot sample a batch of size %d from buffer with size %d." % (n, self.count))
        return self.rng.choice(self.capacity, size=n, p=softmax_probs(self.l))

    def sample(self, n):
        return self.stack_transitions(self.sample_indices(n))

    def batch_of_ids_transitions_and_logits(self, n):
        if self.count < n:
            raise BufferError("Cannot sample a batch of size %d from buffer with size %d." % (n, self.count))
        idx = self.rng.choice(self.count, size=n, replace=False)
        return idx, self.stack_transitions(idx), self.l[np.asarray(idx)]

    def update_priorities(self, idx, values):
        self.l[np.asarray(idx)] = values


class PairAccumulator:
    """Pairs consecutive timesteps into transitions (replay.py:1169-1197)."""

    def __init__(self):
        self.prev = None
        self.prev_action = None

    def reset(self):
        self.prev = None
        self.prev_action = None

    def step(self, timestep, action):
        if timestep.first():
            self.reset()
        out = []
        if self.prev is None:
            if not timestep.first():
                raise ValueError("Expected FIRST timestep, got %s." % str(timestep))
        else:
            out.append(Transition(self.prev.observation, self.prev_action, timestep.reward,
                                  timestep.discount, timestep.observation))
        self.prev, self.prev_action = timestep, action
        return out


# ---- prioritized replay (replay.py:336-784, 1046-1160), restated in array form -------------------------
def is_weights(probs, uniform_prob, exponent, normalize):
    """(uniform_prob / p) ** exponent, optionally scaled so the largest weight is 1 (replay.py:344-376)."""
    w = (uniform_prob / probs) ** exponent
    return w / w.max() if normalize else w


class PrefixTree:
    """Complete binary tree of float64 partial sums over `width` leaves (width = next power of two of the
    capacity): node k = node 2k + node 2k+1, root at 1, leaves from `width` (replay.py:386-396 layout).  The
    additions happen in tree order, which is what makes query results reproducible bit for bit."""

    def __init__(self, capacity):
        self.width = 1
        while self.width < capacity:
            self.width *= 2
        self.node = np.zeros(2 * self.width, dtype=np.float64)

    @property
    def total(self):
        return self.node[1]

    def assign(self, leaves, values):
        for leaf, v in zip(leaves, values):          # in order: a later duplicate overrides an earlier one
            k = self.width + int(leaf)
            self.node[k] = v
            k //= 2
            while k >= 1:
                self.node[k] = self.node[2 * k] + self.node[2 * k + 1]
                k //= 2

    def locate(self, target):
        """Smallest leaf whose inclusive running sum exceeds `target` (replay.py:536-559)."""
        if not 0.0 <= target < self.total:
            raise ValueError("Require 0 <= target < total sum.")
        k = 1
        while k < self.width:
            left = self.node[2 * k]
            if target < left:
                k = 2 * k
            else:
                target -= left
                k = 2 * k + 1
        return k - self.width


class PrioritizedReplay:
    """FIFO items + proportional prioritisation (replay.py:562-784 + 1046-1160).  Leaf allocation mirrors the
    reference: free leaves are a stack initialised 0..capacity-1 and popped from the END, a removed id's leaf is
    pushed back; the `active` list (uniform part of the mixture) uses swap-with-last removal."""

    def __init__(self, capacity, structure, priority_exponent, importance_sampling_exponent,
                 uniform_sample_probability, normalize_weights, random_state):
        self.capacity, self.structure, self.rs = capacity, structure, random_state
        self.alpha, self.beta_fn = priority_exponent, importance_sampling_exponent
        self.usp, self.normalize = uniform_sample_probability, normalize_weights
        self.tree = PrefixTree(capacity)
        self.free = list(range(capacity))
        self.active, self.where = [], {}          # active leaves; leaf -> position in `active`
        self.leaf_of, self.id_of = {}, {}
        self.items = collections.OrderedDict()
        self.t = 0

    @property
    def size(self):
        return len(self.items)

    def _pow(self, p):
        p = np.asarray(p)
        return np.where(p == 0.0, 0.0, p ** self.alpha)

    def add(self, item, priority):
        if len(self.items) == self.capacity:
            old, _ = self.items.popitem(last=False)
            leaf = self.leaf_of.pop(old)
            del self.id_of[leaf]
            pos = self.where.pop(leaf)
            tail = self.active.pop()
            if tail != leaf:
                self.active[pos] = tail
                self.where[tail] = pos
            self.free.append(leaf)
            self.tree.assign([leaf], [0.0])
        leaf = self.free.pop()
        self.where[leaf] = len(self.active)
        self.active.append(leaf)
        self.leaf_of[self.t], self.id_of[leaf] = leaf, self.t
        self.tree.assign([leaf], self._pow([priority]))
        self.items[self.t] = item
        self.t += 1

    def sample_ids(self, n):
        """-> (ids, importance weights); RandomState consumption as replay.py:681-717."""
        draw = self.rs.integers if isinstance(self.rs, np.random.Generator) else self.rs.randint   # replay_circular.py:1017
        uniform_leaves = np.array([self.active[j] for j in draw(self.size, size=n)])
        total = self.tree.total
        if total == 0.0:
            picked = uniform_leaves
        else:
            picked = np.array([self.tree.locate(x) for x in self.rs.uniform(size=n) * total])
        leaves = np.where(self.rs.uniform(size=n) < self.usp, uniform_leaves, picked)
        u = np.asarray(1.0 / self.size)
        pri = self.tree.node[self.tree.width + leaves]
        p = np.full_like(pri, u) if total == 0.0 else pri / total
        probs = (1.0 - self.usp) * p + self.usp * u
        ids = np.array([self.id_of[int(l)] for l in leaves], dtype=np.int64)
        return ids, is_weights(probs, 1.0 / self.size, self.beta_fn(self.t), self.normalize)

    def sample(self, n):
        ids, w = self.sample_ids(n)
        return _stack_fields(self.structure, [self.items[int(i)] for i in ids]), ids, w

    def update_priorities(self, ids, priorities):
        self.tree.assign([self.leaf_of[int(i)] for i in ids], self._pow(np.asarray(priorities)))
