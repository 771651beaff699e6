"""TF checkpoint (train directory or bundle prefix) -> model.pgenw, without TensorFlow.
Usage: python tools/convert_checkpoint.py <log_root>/<exp_name>/train [out.pgenw]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import tf_bundle, weights_io

src = sys.argv[1]
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(src if os.path.isdir(src) else os.path.dirname(src), weights_io.DEFAULT_NAME)
names = tf_bundle.convert_checkpoint(src, out)
print("%d variables -> %s" % (len(names), out))
