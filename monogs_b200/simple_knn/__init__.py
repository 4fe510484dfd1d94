"""Drop-in replacement for the `simple_knn` package (submodules/simple-knn): `simple_knn._C.distCUDA2`."""
