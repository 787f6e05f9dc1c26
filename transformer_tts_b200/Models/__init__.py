"""Drop-in replacements for the FastSpeech2 hot-path classes of the reference's ``Models`` package.

Module names mirror the reference (``Models.fastspeech2``, ``Models.encoder``, ``Models.layers``,
``Models.modules``, ``Models.varianceadaptor``, ``Models.postnets``) so that
``train_fastspeech2.py:29-30`` / ``test_fastspeech2.py:29`` style imports keep working when this
directory is put first on ``sys.path``.
"""
