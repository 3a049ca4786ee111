"""Drop-in `simple_knn` package (the reference ships only the `_C` extension module)."""
