"""Bit-packing helpers of the 1-bit communication back ends (reference: ``deepspeed/runtime/compression``)."""
