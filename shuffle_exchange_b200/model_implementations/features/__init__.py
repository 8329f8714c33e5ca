"""Optional model features."""
