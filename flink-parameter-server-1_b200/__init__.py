"""B200-native asynchronous parameter server (capabilities of flink-parameter-server).

Two engines behind one API:
  * host engine  - arbitrary Python logics / parameter types, exact reference message semantics
                   (``parameter_server.transform*``); the oracle.
  * device engine - HBM-sharded fp32 tables + fused pull-compute-push sm_100a kernels
                   (``parallel.device_engine``, ``ops``, ``models.*.device``).
"""
from .api import (ParameterServer, ParameterServerClient, ParameterServerLogic, PSReceiver,
                  PSSender, RuntimeContext, WorkerLogic, WorkerReceiver, WorkerSender)
from .entities import (EOF, Either, Left, PSToWorker, Pull, PullAnswer, Push, Right, WorkerToPS,
                       last_write_wins, lefts, rights)
from .parameter_server import (default_param_partitioner, default_w_in_partition, hash_func,
                               transform, transform_generic, transform_with_double_model_load,
                               transform_with_model_load)
from .worker_logic import (PullAnswerFuture, WorkerLogicWithFuture, add_blocking_pull_limiter,
                           add_pull_limiter)

__version__ = "0.1.0"
