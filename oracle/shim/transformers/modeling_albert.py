"""ORACLE ONLY. The reference star-imports this module for its namespace
side effects (/root/reference/src/image_model.py:5); only these names are used."""
import math  # noqa: F401
import os  # noqa: F401

import torch  # noqa: F401
import torch.nn as nn  # noqa: F401
import torch.nn.functional as F  # noqa: F401
