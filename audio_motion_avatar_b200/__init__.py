"""B200-native render path of liubingqi7/audio-motion-avatar (src/models/renderer.py)."""
__version__ = "0.1.0"
