"""Arctic Long Sequence Training (ALST) / Ulysses sequence parallelism for HF-style models."""
from .parallel_state_sp import (get_sequence_data_parallel_group, get_sequence_data_parallel_rank,
                                get_sequence_data_parallel_world_size, get_sequence_parallel_group,
                                get_sequence_parallel_rank, get_sequence_parallel_world_size,
                                initialize_sequence_parallel)
from .ulysses_sp import (SequenceTiledCompute, TiledFusedLogitsLoss, TiledLoss, TiledMLP, UlyssesSPAttentionHF,
                         UlyssesSPDataLoaderAdapter, UlyssesSPFwdLossBwdWithLogits, sequence_tiled_compute,
                         sxb_flash_core_attention)
