# Executed as the body of the `int_act_b200` package (see int_act_b200/__init__.py).
__version__ = "0.1.0"
