import threading

import numpy as np
import pytest

from distributedvolunteercomputing_b200 import protocol as proto
from distributedvolunteercomputing_b200.config import Config


def test_control_grammar_and_precedence():
    assert proto.make_control("join", "10.0.0.2:5554") == b"join||10.0.0.2:5554"
    assert proto.parse_control(b"join||a:1") == ("join", "a:1")
    assert proto.parse_control(b"request||a:1") == ("request", "a:1")
    assert proto.parse_control(b"stop||a:1") == ("stop", "a:1")
    assert proto.parse_control(b"end||a:1") == ("end", "a:1")
    # substring matching in the reference's order (server.py:104,126,136,141)
    assert proto.parse_control(b"request||joiner:1")[0] == "join"
    assert proto.parse_control(b"hello")[0] is None


def test_header_roundtrip_and_chunk_pack():
    h = proto.Header("127.0.0.1:5554", "request", [1, 2, 3], 48, 64)
    assert h.encode() == "127.0.0.1:5554||request||1-2-3||48||64"
    assert proto.parse_header(h.encode()) == h
    assert h.tag(2) == "127.0.0.1:5554||request||2"
    rng = np.random.default_rng(0)
    frames = [rng.integers(0, 256, (48, 64, 3), dtype=np.uint8) for _ in range(3)]
    img = proto.pack_chunk(frames)
    assert img.shape == (48, 192, 3)
    back = proto.unpack_chunk(img, 3, 48, 64)
    assert all(np.array_equal(a, b) for a, b in zip(frames, back))


def test_imagezmq_wire_roundtrip_and_timeout():
    pytest.importorskip("zmq")
    port = proto.free_tcp_port()
    hub = proto.ImageHub(f"tcp://*:{port}")
    got = {}

    def serve():
        msg, img = hub.recv_image(timeout_s=5)
        got["msg"], got["img"] = msg, np.array(img)
        hub.send_reply(b"OK")

    t = threading.Thread(target=serve)
    t.start()
    s = proto.ImageSender(f"tcp://127.0.0.1:{port}", timeout_s=5)
    arr = np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3)
    assert s.send_image("id||request||1||2||3", arr) == b"OK"
    t.join()
    assert got["msg"] == "id||request||1||2||3" and np.array_equal(got["img"], arr)
    assert hub.recv_image(timeout_s=0.05) is None
    hub.close()
    # dead peer: the reference blocks forever, we time out and stay usable
    dead = proto.ImageSender(f"tcp://127.0.0.1:{proto.free_tcp_port()}", timeout_s=0.2)
    with pytest.raises(TimeoutError):
        dead.send_image("x", arr)
    with pytest.raises(TimeoutError):
        dead.send_image("x", arr)
    dead.close(); s.close()


def test_config_defaults_env_and_argv(monkeypatch):
    c = Config()
    assert (c.control_port, c.my_port, c.number_of_frames_in_chunk, c.max_buffer_frames, c.max_buffer_chunks) == (9999, 5554, 100, 4000, 40)
    assert c.confidence == 0.2 and c.consider == ("person",) and c.out_width == 400 and c.writer_fps == 30
    assert c.mean == (127.5, 0.0, 0.0) and c.scale == 0.007843
    monkeypatch.setenv("DVC_MY_PORT", "6001")
    monkeypatch.setenv("DVC_MEAN", "127.5,127.5,127.5")
    c = Config.from_argv(["10.0.0.1", "10.0.0.2", "--number-of-frames-in-chunk", "16", "--verbose"], ("server_ip", "own_ip"))
    assert (c.server_ip, c.own_ip, c.my_port, c.number_of_frames_in_chunk, c.verbose) == ("10.0.0.1", "10.0.0.2", 6001, 16, True)
    assert c.mean == (127.5, 127.5, 127.5) and c.addr == "10.0.0.2:6001"
    assert Config.from_argv([], ("server_ip",)).server_ip == "localhost"      # positional-only invocation keeps working


def test_config_rejects_settings_that_would_be_silently_ignored():
    import pytest
    from distributedvolunteercomputing_b200.config import Config
    assert Config().validate().transport == "req_rep"
    with pytest.raises(ValueError, match="PUB/SUB"):
        Config(transport="pub_sub").validate()
    with pytest.raises(ValueError, match="dispatch_policy"):
        Config(dispatch_policy="random").validate()
    with pytest.raises(ValueError):
        Config(number_of_frames_in_chunk=0).validate()
