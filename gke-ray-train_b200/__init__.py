"""Directory marker: the importable host package is `b200_train` inside this directory
(`sys.path.insert(0, "<repo>/gke-ray-train_b200")`; the directory name itself is not a Python identifier)."""
