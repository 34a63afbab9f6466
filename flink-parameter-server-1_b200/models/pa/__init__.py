from .algorithm import (PassiveAggressiveAlgorithm, PassiveAggressiveBinaryAlgorithm,
                        PassiveAggressiveBinaryAlgorithmI, PassiveAggressiveBinaryAlgorithmII,
                        PassiveAggressiveCostBased, PassiveAggressiveFilter, PassiveAggressiveFilterI,
                        PassiveAggressiveFilterII, PassiveAggressiveMultiFilter,
                        PassiveAggressiveMulticlassAlgorithm, PassiveAggressiveOneVersusAll,
                        PassiveAggressiveOneVersusAllI, PassiveAggressiveOneVersusAllII,
                        RandomModelInitializer, binary_initializer, multiclass_initializer,
                        zero_one_cost)
from .parameter_server import (PAWorkerLogic, hash_partitioner_ps, range_partitioner_ps,
                               transform_binary, transform_multiclass,
                               transform_multiclass_with_long_id)
from .sparse import EOFSign, LegacySparseVector, SparseVector
from .offline import (PAOfflineWorker, pa_binary_classification_offline,
                      pa_multi_classification_offline)
