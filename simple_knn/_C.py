"""`simple_knn._C` — distCUDA2 (reference simple-knn/ext.cpp:15-16), backed by das3r_b200."""
from das3r_b200 import load_native as _load

distCUDA2 = _load().distCUDA2
