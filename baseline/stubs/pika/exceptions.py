class AMQPConnectionError(Exception):
    pass
