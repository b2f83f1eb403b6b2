"""Drop-in `utils` package: the reference's hot-path modules (fusion_model, lidar, losses, camera,
evaluation) re-implemented on libtdssf.so behind the same import names.  Put this directory's
parent first on sys.path and the reference's main_*.py / utils/train.py / utils/test.py import it
unchanged."""
