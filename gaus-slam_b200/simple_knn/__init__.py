"""B200-native drop-in for the `simple_knn` package (KNN/ext.cpp:15-17): `from simple_knn._C import distCUDA2`
as scene/Gaussians.py:8 imports it."""
