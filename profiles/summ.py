import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("img/s", round(d["images_per_s"]), "e2e", round(d["e2e"]["images_per_s"]), d["device_ms_per_image"], "xf_frac %.3f"%d["roofline_transform"]["frac"], "ent_ms %.4f"%d["roofline_entropy"]["ms_per_launch"], "launches", d["gpu_launches"], d.get("cpu_baseline"))
