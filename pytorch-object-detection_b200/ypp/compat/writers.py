"""The on-disk detection formats directly downstream of the path.

    pred-file text rows   make_pred_file_yolov3.py:73-86  (make_pred_file_yolov2.py twin)
    COCO results rows     yolo2coco_pred_file_yolov3.py:83-100 (yolo2coco_pred_file_yolov2.py twin)

The reference scripts loop over the kept boxes of ONE image in Python, doing 0-dim tensor
arithmetic and int() per box.  Here the integer columns of every row of a whole batch come from
one kernel (ypp_detection_ints); only the string / dict formatting stays on the host.  Text and
JSON are byte-identical to what the reference loops produce for the same rows.
"""
import json
import os

import torch

from .. import engine


def _host(rows, ints):
    return rows.detach().cpu().numpy(), ints.cpu().numpy()


def pred_file_lines(boxes, class_name_list):
    """boxes [K,6] (one image, what DecodeYoloV3/V2 return) -> the lines the reference writes:
    f'{class_name} {confidence_score} {xmin} {ymin} {xmax} {ymax}\\n'."""
    if boxes.shape[0] == 0:
        return []
    r, v = _host(boxes, engine.detection_ints(boxes, "pred_file"))
    o = boxes.shape[1] - 6
    # format() of a 0-dim tensor is format() of its Python float (torch/_tensor.py:__format__)
    return ["%s %s %d %d %d %d\n" % (class_name_list[int(r[k, o + 5])], format(float(r[k, o + 4])),
                                     v[k, 0], v[k, 1], v[k, 2], v[k, 3]) for k in range(r.shape[0])]


def write_pred_files(det_rows, counts, class_name_list, save_dir, first_img_num=1):
    """det_rows [sum K,7|6] packed image after image, counts [B] -> one f'{img_num:05d}.txt' per image
    (img_num starts at 1 and is incremented before use, make_pred_file_yolov3.py:58,70-71)."""
    os.makedirs(save_dir, exist_ok=True)
    counts = [int(c) for c in (counts.tolist() if torch.is_tensor(counts) else counts)]
    lines = pred_file_lines(det_rows[: sum(counts)], class_name_list)
    off, paths = 0, []
    for i, c in enumerate(counts):
        path = os.path.join(save_dir, "%05d.txt" % (first_img_num + i))
        with open(path, "w") as f:
            f.writelines(lines[off: off + c])
        off += c
        paths.append(path)
    return paths


def coco_results(det_rows, image_ids, widths, heights, input_size, cats_dict, class_name_list, img_idx_base=0):
    """det_rows [sum K,7] (img column = img_idx_base + position in image_ids) or [K,6] (a single image) ->
    list of {"image_id", "category_id", "bbox": [xmin, ymin, w, h], "score"} in row order."""
    if det_rows.shape[0] == 0:
        return []
    # width / cfg['input_size'] is a Python float; multiplying a fp32 tensor by it rounds it to fp32
    scale = torch.tensor([[w / input_size, h / input_size] for w, h in zip(widths, heights)], dtype=torch.float64).float()
    ints = engine.detection_ints(det_rows, "coco", scale_xy=scale, img_idx_base=img_idx_base)
    r, v = _host(det_rows, ints)
    o = det_rows.shape[1] - 6
    out = []
    for k in range(r.shape[0]):
        img = int(r[k, 0] - img_idx_base) if o else 0
        out.append({"image_id": image_ids[img],
                    "category_id": cats_dict[class_name_list[int(r[k, o + 5])]],
                    "bbox": [int(v[k, 0]), int(v[k, 1]), int(v[k, 2]), int(v[k, 3])],
                    "score": float(r[k, o + 4])})
    return out


def write_coco_results(results, results_json_path):
    with open(results_json_path, "w") as f:
        json.dump(results, f, indent=4)              # yolo2coco_pred_file_yolov3.py:102-103
    return results_json_path
