"""lp <-> hp parameter linkage (reference: deepspeed/utils/mixed_precision_linkage.py).  With FlatGroup
storage the link is created once per group by ``tensor_fragment.link_group_params``; these helpers
keep the reference names."""
from .tensor_fragment import link_group_params


def link_hp_params(lp_param_list, flat_hp_partition, gradient_dict, offload_gradient_dict, use_offload,
                   param_group_index, partition_start, partition_size, dp_group, optimizer=None):
    for p in lp_param_list:
        if getattr(p, "_sxb_link", None) is None and optimizer is not None:
            for fg in optimizer.groups:
                if any(p is q for q in fg.params):
                    link_group_params(fg, optimizer)


def lazy_init_hp_params_optimizer_state(lp_param_list, flat_hp_partition, optimizer_state):
    return None
