"""Hand assembly of AMQP 0-9-1 frames for the decode tests, following the reference's encoders
(chana-mq-base/.../model/Frame.scala:217-226, ValueWriter.scala:46-97, ContentHeaderPropertyWriter.scala:31-53,
method/Basic.scala:155-170)."""
import struct

METHOD, HEADER, BODY, HEARTBEAT = 1, 2, 3, 8


def frame(ftype, channel, payload, end=0xCE):
    return struct.pack(">BHI", ftype, channel, len(payload)) + payload + bytes([end])


def shortstr(s):
    b = s if isinstance(s, bytes) else s.encode()
    return bytes([len(b)]) + b


def longstr(b):
    return struct.pack(">I", len(b)) + b


def field(tag, value=None):
    t = tag.encode()
    if tag in "Sx":
        return t + longstr(value)
    if tag == "I":
        return t + struct.pack(">i", value)
    if tag == "D":
        return t + bytes([value[0]]) + struct.pack(">i", value[1])
    if tag == "T":
        return t + struct.pack(">Q", value)
    if tag == "F":
        return t + table(value)
    if tag == "A":
        body = b"".join(field(*v) for v in value)
        return t + struct.pack(">I", len(body)) + body
    if tag == "b":
        return t + struct.pack(">b", value)
    if tag == "d":
        return t + struct.pack(">d", value)
    if tag == "f":
        return t + struct.pack(">f", value)
    if tag == "l":
        return t + struct.pack(">q", value)
    if tag == "s":
        return t + struct.pack(">h", value)
    if tag == "t":
        return t + bytes([1 if value else 0])
    if tag == "V":
        return t
    raise ValueError(tag)


def table(fields):
    """fields: list of (name, tag, value)"""
    body = b"".join(shortstr(n) + field(tag, v) for n, tag, v in fields)
    return struct.pack(">I", len(body)) + body


def publish_method(exchange, key, mandatory=False, immediate=False, channel=1):
    bits = (1 if mandatory else 0) | (2 if immediate else 0)
    return frame(METHOD, channel, struct.pack(">HHH", 60, 40, 0) + shortstr(exchange) + shortstr(key) + bytes([bits]))


def other_method(cls, mth, args=b"", channel=1):
    return frame(METHOD, channel, struct.pack(">HH", cls, mth) + args)


PROP_ORDER = ["contentType", "contentEncoding", "headers", "deliveryMode", "priority", "correlationId", "replyTo",
              "expiration", "messageId", "timestamp", "type", "userId", "appId", "clusterId"]


def content_header(body_size, channel=1, class_id=60, flags_extra=0, **props):
    flags, body = 0, b""
    for i, name in enumerate(PROP_ORDER):
        if name not in props:
            continue
        flags |= 1 << (15 - i)
        v = props[name]
        if name == "headers":
            body += v if isinstance(v, bytes) else table(v)
        elif name in ("deliveryMode", "priority"):
            body += bytes([v])
        elif name == "timestamp":
            body += struct.pack(">Q", v)
        else:
            body += shortstr(v)
    return frame(HEADER, channel, struct.pack(">HHqH", class_id, 0, body_size, flags | flags_extra) + body)


def body_frame(data, channel=1):
    return frame(BODY, channel, data)


def heartbeat():
    return frame(HEARTBEAT, 0, b"")


def message(exchange, key, body=b"", channel=1, frame_max=None, **props):
    out = publish_method(exchange, key, channel=channel) + content_header(len(body), channel=channel, **props)
    step = frame_max or max(len(body), 1)
    for i in range(0, len(body), step):
        out += body_frame(body[i:i + step], channel)
    return out
