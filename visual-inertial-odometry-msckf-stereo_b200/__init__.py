"""B200-native MSCKF back end.  Import as ``msckf_b200`` (see ../msckf_b200/__init__.py)."""
