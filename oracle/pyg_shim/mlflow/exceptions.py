class MlflowException(Exception):
    pass
