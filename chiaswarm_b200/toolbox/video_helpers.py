"""Video helpers (reference: swarm/toolbox/video_helpers.py:6-25 uses moviepy; here OpenCV, which
is what the image ships)."""
from __future__ import annotations

from io import BytesIO

from PIL import Image


def get_frame(mp4_file_path, frame_index=0):
    """JPEG BytesIO of one frame, or None when the file cannot be decoded."""
    try:
        import cv2

        cap = cv2.VideoCapture(str(mp4_file_path))
        try:
            if frame_index:
                cap.set(cv2.CAP_PROP_POS_FRAMES, int(frame_index))
            ok, frame = cap.read()
        finally:
            cap.release()
        if not ok:
            return None
        buf = BytesIO()
        Image.fromarray(cv2.cvtColor(frame, cv2.COLOR_BGR2RGB)).save(buf, format="JPEG")
        return buf
    except Exception as e:
        print(e)
        return None


def frames_to_video(frames, path, codec, fps=8):
    """uint8 HWC RGB frames -> container at `path` (reference: swarm/video/tx2vid.py:79-88)."""
    import cv2

    h, w, _ = frames[0].shape
    writer = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*codec), fps, (w, h))
    if not writer.isOpened():
        raise RuntimeError(f"cannot open video writer for codec {codec}")
    try:
        for frame in frames:
            writer.write(cv2.cvtColor(frame, cv2.COLOR_RGB2BGR))
    finally:
        writer.release()
    return path
