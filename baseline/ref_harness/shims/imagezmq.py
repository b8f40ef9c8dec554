"""Minimal stand-in for the `imagezmq` package (not installed here, no network): same classes,
same ZeroMQ REQ/REP wire format (JSON {msg,dtype,shape} + raw bytes, b'OK' replies) that the
reference relies on (server.py:43,112; worker.py:64,164).  Used only to run the UNMODIFIED
reference for baseline numbers."""
import numpy as np
import zmq


class ImageSender:
    def __init__(self, connect_to="tcp://127.0.0.1:5555", REQ_REP=True):
        self.zmq_context = zmq.Context.instance()
        if REQ_REP:
            self.zmq_socket = self.zmq_context.socket(zmq.REQ)
            self.zmq_socket.connect(connect_to)
        else:
            self.zmq_socket = self.zmq_context.socket(zmq.PUB)
            self.zmq_socket.bind(connect_to)
        self.REQ_REP = REQ_REP

    def send_image(self, msg, image):
        if not image.flags["C_CONTIGUOUS"]:
            image = np.ascontiguousarray(image)
        md = dict(msg=msg, dtype=str(image.dtype), shape=image.shape)
        self.zmq_socket.send_json(md, zmq.SNDMORE)
        self.zmq_socket.send(image, 0, copy=True, track=False)
        if self.REQ_REP:
            return self.zmq_socket.recv()

    def close(self):
        self.zmq_socket.close()


class ImageHub:
    def __init__(self, open_port="tcp://*:5555", REQ_REP=True):
        self.zmq_context = zmq.Context.instance()
        if REQ_REP:
            self.zmq_socket = self.zmq_context.socket(zmq.REP)
            self.zmq_socket.bind(open_port)
        else:
            self.zmq_socket = self.zmq_context.socket(zmq.SUB)
            self.zmq_socket.setsockopt(zmq.SUBSCRIBE, b"")
            self.zmq_socket.connect(open_port)
        self.REQ_REP = REQ_REP

    def recv_image(self, copy=True):
        # copy=True: the reference keeps crops of this array alive across later recv calls and hands
        # them to cv2.VideoWriter from another thread; with zero-copy zmq frames that segfaults here
        # (cv2 4.13 / pyzmq 27), so the shim owns its bytes.  The wire format is unchanged.
        md = self.zmq_socket.recv_json()
        msg = self.zmq_socket.recv(copy=copy, track=False)
        A = np.frombuffer(bytearray(msg) if copy else msg, dtype=md["dtype"])
        return md["msg"], A.reshape(md["shape"])

    def send_reply(self, reply_message=b"OK"):
        self.zmq_socket.send(reply_message)

    def close(self):
        self.zmq_socket.close()
