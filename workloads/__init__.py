"""Seeded synthetic inputs (frames, YOLO heads, detections) shared by the tests, bench.py and
the examples.  Pure data generation: no reference arithmetic lives here."""
