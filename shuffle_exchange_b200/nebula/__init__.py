"""Nebula asynchronous checkpoint service configuration (reference: ``deepspeed/nebula``)."""
