"""Dropout RNG state tracker — host-side mirror of
paddle.distributed.fleet.meta_parallel.get_rng_state_tracker() as the reference uses it:

  ppfleetx/utils/env.py:27-46      tracker.add('global_seed', seed + data_world_rank)
                                   tracker.add('local_seed', seed + 123 + mp_rank*10 + pp_rank*1000)
  hybrid_model.py:280              with get_rng_state_tracker().rng_state('local_seed'):  attention dropout
  hybrid_model.py:546, :556        with get_rng_state_tracker().rng_state('global_seed'): dropout1 / dropout2

A named state here is (seed, offset). The kernels derive the keep mask as a pure function of
(seed, offset, element index) (csrc/philox.cuh), so "consuming" randomness is just taking the
current offset and advancing it by one per dropout call. Saving / restoring states therefore
reproduces masks exactly, which is what the reference's recompute wrappers
(hybrid_model.py:332-335, 406-408, 537-539) need.
"""
from __future__ import annotations

import contextlib
from typing import Dict, Tuple

MODEL_PARALLEL_RNG = "model_parallel_rng"


class RNGStatesTracker:
    def __init__(self):
        self.states_: Dict[str, list] = {}
        self.seeds_ = set()
        self._current: str | None = None

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()
        self._current = None

    def add(self, name: str, seed: int):
        # same error behaviour as Paddle's tracker: duplicate seed or name is an error
        if seed in self.seeds_:
            raise ValueError(f"seed {seed} already exists")
        if name in self.states_:
            raise ValueError(f"state {name} already exists")
        self.seeds_.add(seed)
        self.states_[name] = [int(seed), 0]

    def get_states_tracker(self) -> Dict[str, Tuple[int, int]]:
        return {k: (v[0], v[1]) for k, v in self.states_.items()}

    def set_states_tracker(self, states: Dict[str, Tuple[int, int]]):
        self.states_ = {k: [int(v[0]), int(v[1])] for k, v in states.items()}

    @contextlib.contextmanager
    def rng_state(self, name: str = MODEL_PARALLEL_RNG):
        if name not in self.states_:
            raise ValueError(f"state {name} does not exist")
        prev = self._current
        self._current = name
        try:
            yield
        finally:
            self._current = prev

    def draw(self, name: str | None = None) -> Tuple[int, int]:
        """(seed, offset) for one dropout call on stream `name` (default: the active context)."""
        name = name or self._current
        if name is None or name not in self.states_:
            raise ValueError(f"no RNG state {name!r}; call set_seed() / tracker.add() first")
        st = self.states_[name]
        seed, offset = st[0], st[1]
        st[1] += 1
        return seed, offset


_TRACKER = RNGStatesTracker()


def get_rng_state_tracker() -> RNGStatesTracker:
    return _TRACKER


def set_seed(seed: int, mp_rank: int = 0, pp_rank: int = 0, data_world_rank: int = 0) -> None:
    """ppfleetx/utils/env.py:27-46 (the two dropout streams only)."""
    tracker = get_rng_state_tracker()
    tracker.reset()
    tracker.add("global_seed", seed + data_world_rank)
    tracker.add("local_seed", seed + 123 + mp_rank * 10 + pp_rank * 1000)
