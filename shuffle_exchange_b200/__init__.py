"""shuffle_exchange_b200 — a B200-native (sm_100a) training engine with the capabilities of the
vituslcx/Shuffle-exchange DeepSpeed fork.  See DESIGN.md / SURVEY.md."""
__version__ = "0.1.0"
