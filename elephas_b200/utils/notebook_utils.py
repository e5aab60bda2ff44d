"""Notebook detection (reference: ``elephas/utils/notebook_utils.py:1-9``)."""


def is_running_in_notebook() -> bool:
    """True when executed inside an IPython *kernel* (Jupyter), False in scripts and plain shells."""
    try:
        shell = get_ipython()  # noqa: F821 - injected by IPython
    except NameError:
        return False
    return 'IPKernelApp' in getattr(shell, 'config', {})
