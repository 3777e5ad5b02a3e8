import base64
import io

from PIL import Image


def png_data_uri(size=(96, 64), color=(200, 30, 30)) -> str:
    buf = io.BytesIO()
    Image.new("RGB", size, color).save(buf, format="PNG")
    return "data:image/png;base64," + base64.b64encode(buf.getvalue()).decode()


def decode_artifact_image(artifact) -> Image.Image:
    return Image.open(io.BytesIO(base64.b64decode(artifact["blob"])))
