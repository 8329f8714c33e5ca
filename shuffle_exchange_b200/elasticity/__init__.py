from .config import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize
from .elasticity import (compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config,
                         get_candidate_batch_sizes, get_valid_gpus)
from .utils import is_torch_elastic_compatible
