"""Utilities: flat serialisation, gloo messaging, CSV logging, checkpoints, device timing."""
