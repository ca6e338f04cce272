"""In-memory stand-in for the ``pika`` AMQP client (the real one and a RabbitMQ broker are not
installable on this image).  It implements only what the reference touches
(``server.py:102-108,178-179,194-203``; ``src/RpcClient.py:42-56,174-188``): a process-global
dict of FIFO queues behind ``BlockingConnection().channel()``.  Nothing here is product code; it
exists so the UNMODIFIED reference can be driven in one process as an oracle / baseline."""
from collections import defaultdict, deque

from . import exceptions  # noqa: F401

QUEUES = defaultdict(deque)
CONSUMERS = {}


class PlainCredentials:
    def __init__(self, username, password):
        self.username, self.password = username, password


class ConnectionParameters:
    def __init__(self, host=None, port=None, virtual_host=None, credentials=None, **kw):
        self.host, self.port = host, port


class _Method:
    def __init__(self, tag):
        self.delivery_tag = tag


class _Props:
    reply_to = None


class Channel:
    _tag = 0

    def queue_declare(self, queue, durable=False, **kw):
        QUEUES[queue]
        return None

    def basic_qos(self, prefetch_count=1):
        pass

    def basic_consume(self, queue, on_message_callback, **kw):
        CONSUMERS[queue] = on_message_callback

    def basic_publish(self, exchange, routing_key, body, **kw):
        QUEUES[routing_key].append(body)

    def basic_get(self, queue, auto_ack=True):
        q = QUEUES[queue]
        if q:
            Channel._tag += 1
            return _Method(Channel._tag), _Props(), q.popleft()
        return None, None, None

    def basic_ack(self, delivery_tag=None):
        pass

    def queue_delete(self, queue):
        QUEUES.pop(queue, None)

    def queue_purge(self, queue):
        QUEUES[queue].clear()

    def start_consuming(self):
        raise RuntimeError("pika stub: pump messages with baseline.ref_loader.pump() instead")


class BlockingConnection:
    def __init__(self, params=None):
        self.params = params

    def channel(self):
        return Channel()

    def close(self):
        pass


def reset():
    QUEUES.clear()
    CONSUMERS.clear()
