"""Contact-sheet workflow (reference: swarm/toolbox/stitch.py:10-110). Pure CPU: N result images
are shrunk to <= 144 px labelled thumbnails and pasted row-major into a ceil(sqrt(N))^2 grid; an
HTML image-map description goes into pipeline_config["image_map"]."""
from __future__ import annotations

import io
import math

from PIL import Image, ImageDraw

from .. import net
from ..output_processor import image_to_buffer, make_result, make_thumbnail

thumb_size = 144


def stitch_callback(device_id, model_name, **kwargs):
    print("Stitching...")
    jobs = kwargs["jobs"]
    images = download_images([job["resultUri"] for job in jobs])
    resized = resize_images(images)
    sheet = stitch_images(resized)
    buffer = image_to_buffer(sheet, "image/jpeg", "web_high")
    results = {"primary": make_result(buffer, make_thumbnail(buffer), "image/jpeg")}
    return results, {"model_name": model_name, "image_map": generate_image_map(resized, jobs)}


def download_images(image_urls):
    return [Image.open(io.BytesIO(net.fetch(url))) for url in image_urls]


def resize_images(images, size=(thumb_size, thumb_size)):
    out = []
    for index, image in enumerate(images):
        width, height = image.size
        aspect = float(width) / float(height)
        if width > height:
            new_w = min(size[0], width)
            new_h = int(new_w / aspect)
        else:
            new_h = min(size[1], height)
            new_w = int(new_h * aspect)
        new_w = max(1, min(new_w, size[0]))
        new_h = max(1, min(new_h, size[1]))
        thumb = image.convert("RGB").resize((new_w, new_h), Image.Resampling.LANCZOS)
        ImageDraw.Draw(thumb).text((10, 10), str(index + 1), fill=(255, 255, 255))
        out.append(thumb)
    return out


def _cells(count):
    """Top-left corner of every thumbnail cell, row-major."""
    per_row = math.ceil(math.sqrt(count)) if count else 0
    return [((i % per_row) * thumb_size, (i // per_row) * thumb_size) for i in range(count)], per_row


def stitch_images(resized_images):
    cells, per_row = _cells(len(resized_images))
    side = thumb_size * per_row
    sheet = Image.new("RGB", (side, side))
    for image, corner in zip(resized_images, cells):
        sheet.paste(image, corner)
    return sheet


def generate_image_map(resized_images, jobs):
    cells, _ = _cells(len(resized_images))
    areas = []
    for index, (x, y) in enumerate(cells):
        job = jobs[index]
        href = job["resultUri"]
        areas.append({
            "shape": "rect",
            "coords": f"{x},{y},{x + thumb_size},{y + thumb_size}",
            "href": href,
            "alt": job.pop("model_name", f"Image {index + 1}"),
            "filename": job.pop("fileName", href),
        })
    return areas
