"""Drop-in package name for DAS3R's `from simple_knn._C import distCUDA2` (scene/gaussian_model.py:21)."""
