"""Matrix-factorisation entry points on the host engine.

Parity:
  * ``ps_offline_mf``                     - ``PSOfflineMatrixFactorization.scala:45-106``
  * ``ps_online_mf``                      - ``PSOnlineMatrixFactorization.scala:39-75`` (arguments are
    passed by name, so ``user_memory`` / ``negative_sample_rate`` are not swapped - SURVEY Q1)
  * ``ps_top_k_generator``                - ``PSTopKGenerator.scala:47-107``
  * ``ps_online_learner_and_generator``   - ``PSOnlineMatrixFactorizationAndTopKGenerator.scala:50-100``
  * ``OnlineFactorModelBuilder``          - ``OnlineFactorModelBuilder.scala:5-12``

The device twins with the same signatures live in ``models/mf/device.py``.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import numpy as np

from ...entities import EOF, Left, Right
from ...ops import philox
from ...parallel.streams import DataStream, as_stream
from ...parameter_server import transform, transform_with_double_model_load
from ...server import SimplePSLogic
from ...utils.eof import WrapWithEOF, flat_map_with_eof
from ...worker_logic import add_pull_limiter
from .collect_topk import CollectTopKFromEachWorker
from .factors import (FactorInitializerDescriptor, FactorUpdater,
                      RangedRandomFactorInitializerDescriptor, SGDUpdater)
from .pruning import COORD, LI, LEMPPruningStrategy
from .rating import IDGenerator, Rating, hash_partition
from .vector import attach_length, vector_sum
from .workers import (PSOfflineMatrixFactorizationWorker, PSOnlineMatrixFactorizationAndTopKGeneratorWorker,
                      PSOnlineMatrixFactorizationWorker, PSTopKGeneratorWorker)


def ps_offline_mf(src, num_factors: int = 10, range_min: float = -0.01, range_max: float = 0.01,
                  learning_rate: float = 0.01, negative_sample_rate: int = 0, user_memory: int = 128,
                  iterations: int = 1, pull_limit: int = 1600, worker_parallelism: int = 1,
                  ps_parallelism: int = 1, iteration_wait_time: int = 10000, *, seed: int = 0,
                  shuffle: bool = True, lambda_: float = 0.0, engine: str = "host", **engine_kw):
    """Returns the stream of ``Left((user, vec))`` / ``Right((item, vec))`` updates.
    ``engine="device"``: the same job on the GPU engine (``models/mf/device.py``; ``ps_parallelism``
    becomes the number of item shards on one GPU, one shard per rank under torchrun)."""
    if _use_device(engine):
        from .device import ps_offline_mf_device
        return ps_offline_mf_device(src, num_factors, range_min, range_max, learning_rate, negative_sample_rate,
                                    user_memory, iterations, pull_limit, seed=seed, shuffle=shuffle,
                                    lambda_=lambda_, n_shards=_device_shards(ps_parallelism), **engine_kw)
    ratings = flat_map_with_eof(src, WrapWithEOF, worker_parallelism,
                                lambda key, n: key % n, lambda r: r.user)
    init_desc = RangedRandomFactorInitializerDescriptor(num_factors, range_min, range_max, seed,
                                                        philox.STREAM_ITEM)
    worker = PSOfflineMatrixFactorizationWorker(
        num_factors, range_min, range_max, learning_rate, negative_sample_rate, user_memory,
        iterations, seed=seed, shuffle=shuffle, lambda_=lambda_)
    logic = add_pull_limiter(worker, pull_limit)
    init = init_desc.open()
    return transform(ratings, logic, lambda i: init.next_factor(i), vector_sum,
                     worker_parallelism, ps_parallelism, iteration_wait_time, **engine_kw)


def ps_online_mf(src, num_factors: int = 10, range_min: float = -0.01, range_max: float = 0.01,
                 learning_rate: float = 0.01, negative_sample_rate: int = 9, user_memory: int = 128,
                 pull_limit: int = 1600, worker_parallelism: int = 1, ps_parallelism: int = 1,
                 iteration_wait_time: int = 10000, *, seed: int = 0, lambda_: float = 0.0,
                 engine: str = "host", **engine_kw):
    """``engine="device"``: micro-batched launches (or ``streaming=True``: the persistent ring consumer)
    with the device negative sampler honouring ``user_memory`` (``models/mf/device.py``)."""
    if _use_device(engine):
        from .device import ps_online_mf_device
        return ps_online_mf_device(src, num_factors, range_min, range_max, learning_rate, negative_sample_rate,
                                   user_memory, pull_limit, seed=seed, **engine_kw)
    init = RangedRandomFactorInitializerDescriptor(num_factors, range_min, range_max, seed,
                                                   philox.STREAM_ITEM).open()
    worker = PSOnlineMatrixFactorizationWorker(
        num_factors, range_min, range_max, learning_rate,
        user_memory=user_memory, negative_sample_rate=negative_sample_rate, seed=seed, lambda_=lambda_)
    logic = add_pull_limiter(worker, pull_limit)
    part = as_stream(src, 1).partition_custom(lambda r: r.user, worker_parallelism)
    return transform(part, logic, ps_logic=SimplePSLogic(lambda i: init.next_factor(i), vector_sum),
                     worker_parallelism=worker_parallelism, ps_parallelism=ps_parallelism,
                     iteration_wait_time=iteration_wait_time, **engine_kw)


def _use_device(engine: str) -> bool:
    if engine not in ("host", "device", "auto"):
        raise ValueError("engine must be 'host', 'device' or 'auto'")
    if engine == "host":
        return False
    from ...parallel import device_engine
    ok, why = device_engine.device_ready()
    if not ok:
        device_engine.fallback(why, strict=False)
    return ok


def _device_shards(ps_parallelism: int):
    from ...parallel.device_context import get_context
    return None if get_context().distributed else max(1, min(int(ps_parallelism), 16))


def device_pruning(strategy: LEMPPruningStrategy, candidate_filters: bool = False) -> bool:
    """What the device top-K does with a ``LEMPPruningStrategy``
    (``pruning/LEMPPruningStrategy.scala:6-50``).  The default kernel scores whole item tiles on the tensor
    cores and applies the LENGTH bound per bucket (skip every later tile once ``max|p| * |q| <= theta``).
    COORD and INCR are *per-candidate* filters that only avoid scalar dot products of items that cannot
    enter the top-K - they never change the result, and a tile's MMA is issued for all its items anyway.
    So every strategy returns exactly what the reference's would; for the ones with a COORD / INCR part a
    ``DeviceFallbackWarning``-class note says that this part is not executed on that path.
    ``candidate_filters=True`` (an ``engine="device"`` keyword of the front-ends) runs the reference's own
    bucket loop with the chosen strategy's filters on CUDA cores instead (``csrc/lemp_filters.cu``,
    ``ops.topk.lemp_topk_filtered``): same results, the reference's pruned / scored counts, not the fast path."""
    from .pruning import LENGTH
    if not candidate_filters and not isinstance(strategy, LENGTH):
        import warnings
        from ...parallel.device_engine import DeviceFallbackWarning
        warnings.warn(f"engine='device': pruning strategy {strategy!r} runs as the LENGTH bucket bound; its "
                      "COORD / INCR candidate filters do not change results and are not executed on the "
                      "tensor-core path (candidate_filters=True runs them on CUDA cores)", DeviceFallbackWarning,
                      stacklevel=3)
    return True


def _broadcast_queries(src, worker_parallelism: int) -> DataStream:
    """Each query goes to every worker under one rating id (X7)."""
    def enrich(r: Rating):
        rid = IDGenerator.next()
        return [r.enrich(i, rid) for i in range(worker_parallelism)]
    return as_stream(src, 1).flat_map(enrich).partition_custom(lambda x: x.target_worker, worker_parallelism)


def ps_top_k_generator(src, model, num_factors: int = 10, user_memory: int = 0, K: int = 100,
                       worker_k: int = 75, bucket_size: int = 100,
                       pruning_algorithm: LEMPPruningStrategy = COORD(), pull_limit: int = 1600,
                       worker_parallelism: int = 4, ps_parallelism: int = 4,
                       iteration_wait_time: int = 10000, *, engine: str = "host", **engine_kw):
    """``engine="device"``: tcgen05 scoring of the local item tiles + cross-GPU merge
    (``models/mf/device_topk.py``); see ``device_pruning`` for how ``pruning_algorithm`` is honoured.

    ``model`` yields ``Left((user, (len, vec)))`` -> PS and ``Right((item, (len, vec)))`` -> the
    worker ``item % W`` (placement by key).  Unknown users answer an empty top-K.
    Returns ``[(item, timestamp, [(score, item), ...]), ...]``."""
    if _use_device(engine):
        from .device_topk import ps_top_k_generator_device
        filters = bool(engine_kw.pop("candidate_filters", False))
        return ps_top_k_generator_device(src, model, num_factors, user_memory, K, worker_k,
                                         use_pruning=device_pruning(pruning_algorithm, filters),
                                         candidate_filters=pruning_algorithm if filters else None,
                                         bucket_size=bucket_size, **engine_kw)
    invalid = (-1, np.zeros(0))
    worker = PSTopKGeneratorWorker(worker_k, bucket_size, worker_parallelism, pruning_algorithm)
    logic = add_pull_limiter(worker, pull_limit)
    ps_logic = SimplePSLogic(lambda _i: invalid, lambda _old, new: new)
    out = transform_with_double_model_load(model, worker_owner=lambda i, w: i % w)(
        _broadcast_queries(src, worker_parallelism), logic, ps_logic,
        lambda m: hash_partition(m.param_id, ps_parallelism), None,
        worker_parallelism, ps_parallelism, iteration_wait_time, **engine_kw)
    merged = CollectTopKFromEachWorker(K, user_memory, worker_parallelism).process(out)
    return [(item, ts, topk) for (_user, item, ts, topk) in merged]


def ps_online_learner_and_generator(src, num_factors: int = 10, range_min: float = -0.001,
                                    range_max: float = 0.001, learning_rate: float = 0.01,
                                    negative_sample_rate: int = 0, user_memory: int = 65535,
                                    K: int = 100, worker_k: int = 75, bucket_size: int = 100,
                                    pruning_algorithm: LEMPPruningStrategy = LI(5, 2.5),
                                    pull_limit: int = 500, worker_parallelism: int = 4,
                                    ps_parallelism: int = 4, iteration_wait_time: int = 10000, *,
                                    seed: int = 0, engine: str = "host", **engine_kw):
    """Prequential recommend-then-learn.  Returns ``[(user, item, timestamp, topK), ...]``.
    ``engine="device"`` needs ``n_users`` and ``n_items`` (table sizes) among the keyword arguments."""
    if _use_device(engine):
        from .device_topk import ps_online_learner_and_generator_device
        filters = bool(engine_kw.pop("candidate_filters", False))
        out, _lg = ps_online_learner_and_generator_device(
            src, num_factors=num_factors, range_min=range_min, range_max=range_max, learning_rate=learning_rate,
            negative_sample_rate=negative_sample_rate, user_memory=user_memory, K=K, worker_k=worker_k,
            pull_limit=pull_limit, seed=seed, use_pruning=device_pruning(pruning_algorithm, filters),
            candidate_filters=pruning_algorithm if filters else None, bucket_size=bucket_size, **engine_kw)
        return out
    item_desc = RangedRandomFactorInitializerDescriptor(num_factors, range_min, range_max, seed,
                                                        philox.STREAM_ITEM)
    user_init = RangedRandomFactorInitializerDescriptor(num_factors, range_min, range_max, seed,
                                                        philox.STREAM_USER).open()
    worker = PSOnlineMatrixFactorizationAndTopKGeneratorWorker(
        negative_sample_rate=negative_sample_rate, user_memory=user_memory, worker_k=worker_k,
        bucket_size=bucket_size, pruning_algorithm=pruning_algorithm,
        worker_parallelism=worker_parallelism, factor_init_desc=item_desc,
        factor_update=SGDUpdater(learning_rate), seed=seed)
    logic = add_pull_limiter(worker, pull_limit)
    ps_logic = SimplePSLogic(lambda u: attach_length(user_init.next_factor(u)),
                             lambda old, delta: attach_length(vector_sum(old[1], delta[1])))
    out = transform(_broadcast_queries(src, worker_parallelism), logic, ps_logic=ps_logic,
                    worker_parallelism=worker_parallelism, ps_parallelism=ps_parallelism,
                    iteration_wait_time=iteration_wait_time, **engine_kw)
    return CollectTopKFromEachWorker(K, user_memory, worker_parallelism).process(out)


class OnlineFactorModelBuilder:
    """Unused upstream (an abstract hook); kept for API parity."""

    def build_model(self, ratings, factor_init: FactorInitializerDescriptor,
                    factor_update: FactorUpdater, parameters: Dict[str, str]):
        raise NotImplementedError
