"""B200-native pointer-generator beam-search decode behind the TFModel pipeline stage.

The directory name carries a hyphen (it mirrors the reference repository's name), so it is
imported through ``pgen_pkg.load()`` at the repo root under the module name
``textsummarization_on_flink_b200``.
"""
__all__ = ["weights_io", "synth"]
