"""Drop-in import name: ``from torchkafka import KafkaDataset, auto_commit``
(the name the reference installs under, ``setup.py:25-30``, ``README.md:38,88``)."""
from torchkafka_b200 import *  # noqa: F401,F403
from torchkafka_b200 import __all__, __version__  # noqa: F401
