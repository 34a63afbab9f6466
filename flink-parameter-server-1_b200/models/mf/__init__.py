from .collect_topk import CollectTopKFromEachWorker
from .factors import (FactorInitializer, FactorInitializerDescriptor, FactorUpdater,
                      PseudoRandomFactorInitializerDescriptor, RandomFactorInitializerDescriptor,
                      RangedRandomFactorInitializer, RangedRandomFactorInitializerDescriptor,
                      SGDUpdater)
from .jobs import (OnlineFactorModelBuilder, ps_offline_mf, ps_online_learner_and_generator,
                   ps_online_mf, ps_top_k_generator)
from .pruning import (COORD, INCR, LC, LENGTH, LI, LEMPPruningStrategy, coord_pruning, from_string,
                      incr_pruning, length_pruning)
from .rating import IDGenerator, Rating, RichRating, TopKQueue, hash_partition
from .vector import (FactorIsNotANumberException, attach_length, dot_product, vector_length_sqr,
                     vector_sum)
from .workers import (BaseMFWorkerLogic, PSOfflineMatrixFactorizationWorker,
                      PSOnlineMatrixFactorizationAndTopKGeneratorWorker,
                      PSOnlineMatrixFactorizationWorker, PSTopKGeneratorWorker, lemp_top_k)
