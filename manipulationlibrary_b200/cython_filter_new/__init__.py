"""Drop-in for the reference package cython_filter_new (the reference's only native component)."""
