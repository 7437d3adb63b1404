"""Module shipped with ``raydp-submit --py-files`` (see ``pyfiles_main.py``).

Imported by the driver script and, by reference, inside executor-side functions, which proves
the file reached both the driver's and the workers' import paths.
"""


def scale(value, factor=3):
    return value * factor


def describe(rows):
    rows = list(rows)
    return {"rows": len(rows), "sum": sum(rows)}
