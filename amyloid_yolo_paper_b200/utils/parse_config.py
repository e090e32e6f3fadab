"""Darknet ``.cfg`` / ``.data`` readers.

Host-side mirror of the reference's ``utils/parse_config.py:3-36``: the block
list it returns is what ``models.Darknet`` (and every caller that pokes at
``model.module_defs``) sees, so the quirks are kept on purpose:

* every value stays a *string* (``'filters': '32'``);
* a ``[convolutional]`` block starts life with the *integer* ``0`` under
  ``batch_normalize`` and only becomes the string ``'1'`` when the cfg says so
  (the weight loader relies on the truthiness difference, ``models.py:276``);
* lines starting with ``#`` and empty lines are dropped, whitespace around
  keys/values is trimmed.
"""


def parse_model_config(path):
    """cfg file -> list of dict, first entry is the ``[net]`` block."""
    with open(path, "r") as fh:
        raw = fh.read().split("\n")
    blocks = []
    for line in raw:
        if not line or line.startswith("#"):
            continue
        line = line.strip()
        if line.startswith("["):
            kind = line[1:-1].rstrip()
            block = {"type": kind}
            if kind == "convolutional":
                block["batch_normalize"] = 0
            blocks.append(block)
        else:
            key, value = line.split("=")
            blocks[-1][key.rstrip()] = value.strip()
    return blocks


def parse_data_config(path):
    """``.data`` file -> dict of strings (defaults as in the reference,
    ``utils/parse_config.py:23-36``)."""
    options = {"gpus": "0,1,2,3", "num_workers": "10"}
    with open(path, "r") as fh:
        for line in fh:
            line = line.strip()
            if line == "" or line.startswith("#"):
                continue
            key, value = line.split("=")
            options[key.strip()] = value.strip()
    return options
