"""Importable name of the package that lives in ``aliyun-gpushare-device-plugin_b200/``.

The task fixes the directory name, and a hyphenated directory cannot be imported
directly, so this shim extends its own ``__path__`` with that directory: every
``gpushare_b200.<module>`` resolves to ``aliyun-gpushare-device-plugin_b200/<module>.py``.
"""
import os as _os

_PKG_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                         "aliyun-gpushare-device-plugin_b200")
__path__.append(_PKG_DIR)
PACKAGE_DIR = _PKG_DIR
REPO_ROOT = _os.path.dirname(_PKG_DIR)
