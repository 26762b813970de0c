"""Streaming hand-off of the archive to the upload (second half): the consumer side of
``dsg_pack_fd``.

The reference packs the whole archive into a temp file, stats it, and only then uploads, because the remote script
waits for an exact byte count that has to be known before the first byte is sent
(``/root/reference/pkg/devspace/sync/upstream.go:353-376`` stat, ``:387-407`` the remote loop on ``$fileSize``,
``:432`` ``io.Copy``). The compressed size of a stream that is still being produced is unknown, so the size contract
becomes a FRAME contract: the archive travels as frames ``<hex length>\\n<length bytes>``, a frame of length 0 ends
it. The remote side needs nothing beyond what the reference's script already uses (``sh``, ``head -c``, ``tar``);
``read`` consumes the length line byte by byte, ``head -c`` exactly the frame. The local side reads the archive from
the pipe that ``dsg_pack_fd`` writes into, so the first frames leave while the GPU is still compressing the rest.

``upload_archive_streaming`` mirrors ``uploadArchive`` (upstream.go:379-459): same START / END acknowledgements, same
temp-file name, same ``tar xzpf`` into DestPath; ``produce(fd)`` is the packer writing the archive to ``fd``.
Host code only: pipes and a shell, no arithmetic of the pack path.
"""
from __future__ import annotations

import os
import threading
from typing import BinaryIO, Callable, Optional

START_ACK = "START"   # sync_config.go:23-24
END_ACK = "DONE"      # sync_config.go: EndAck
ERROR_ACK = "ERROR"
FRAME_BYTES = 1 << 20


def remote_script(dest_path: str) -> str:
    """What replaces the remote command of upstream.go:387-407: frames in, ``tar xzpf`` at the end-of-archive frame."""
    return ("tmpFile=\"/tmp/devspace-upstream\";\n"
            "mkdir -p /tmp;\n"
            "mkdir -p '" + dest_path + "';\n"
            ": >\"$tmpFile\";\n"
            "echo \"" + START_ACK + "\";\n"
            "while IFS= read -r frameLen; do\n"
            "  n=$(printf '%d' \"0x$frameLen\");\n"
            "  if [ \"$n\" = \"0\" ]; then break; fi;\n"
            "  head -c \"$n\" >>\"$tmpFile\";\n"
            "done;\n"
            "if tar xzpf \"$tmpFile\" -C '" + dest_path + "/.' 2>/tmp/devspace-upstream-error; then echo \"" + END_ACK +
            "\"; else echo \"" + ERROR_ACK + "\"; fi;\n")


def _wait_till(keyword: str, stdout: BinaryIO) -> None:
    """waitTill (upstream.go / util.go): read lines until the acknowledgement arrives."""
    while True:
        line = stdout.readline()
        if not line:
            raise IOError(f"[Upstream] remote side closed the stream while waiting for {keyword}")
        text = line.decode("utf-8", "replace").strip()
        if text == keyword:
            return
        if text == ERROR_ACK:
            raise IOError("[Upstream] remote tar failed")


def send_frames(src_fd: int, stdin: BinaryIO, frame_bytes: int = FRAME_BYTES) -> int:
    """Reads the archive from ``src_fd`` until EOF and writes it to ``stdin`` as frames; returns the archive bytes."""
    total = 0
    while True:
        chunk = os.read(src_fd, frame_bytes)
        if not chunk:
            break
        stdin.write(b"%x\n" % len(chunk))
        stdin.write(chunk)
        total += len(chunk)
    stdin.write(b"0\n")
    stdin.flush()
    return total


def upload_archive_streaming(stdin: BinaryIO, stdout: BinaryIO, dest_path: str, produce: Callable[[int], None],
                             logf: Optional[Callable[[str], None]] = None) -> int:
    """uploadArchive (upstream.go:379-459) without the temp file and without the size: ``produce(fd)`` writes the
    archive into a pipe (dsg_pack_fd) on a thread of its own while this thread frames what has arrived so far and
    sends it. Returns the number of archive bytes sent; raises when the producer or the remote tar fails."""
    stdin.write(remote_script(dest_path).encode())
    stdin.flush()
    _wait_till(START_ACK, stdout)
    r, w = os.pipe()
    failure = []

    def run():
        try:
            produce(w)
        except BaseException as e:  # noqa: BLE001 -- handed to the caller's thread
            failure.append(e)
        finally:
            os.close(w)

    t = threading.Thread(target=run)
    t.start()
    try:
        sent = send_frames(r, stdin)
    finally:
        os.close(r)
        t.join()
    if failure:
        raise failure[0]
    if logf:
        logf(f"[Upstream] Uploaded {sent} bytes in frames")
    _wait_till(END_ACK, stdout)
    return sent
