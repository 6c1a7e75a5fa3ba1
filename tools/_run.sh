timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for r in dense sparse; do python tools/bench_d1.py --regime $r; done
python tools/bench_d1.py --workload yolov3_608_coco_b32 --regime dense
python tools/bench_d1.py --workload yolov4tiny_416_person_b256 --regime dense
