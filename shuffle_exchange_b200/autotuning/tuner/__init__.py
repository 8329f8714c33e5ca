from .index_based_tuner import GridSearchTuner, RandomTuner
from .model_based_tuner import ModelBasedTuner
