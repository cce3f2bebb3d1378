"""atooms_b200 — B200-native pair-interaction / MD backend for atooms."""
__version__ = '0.1.0'
