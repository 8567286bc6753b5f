"""Regenerates the committed fixtures from the reference checkout (run in the build container only;
/root/reference does not exist on the GPU box).

  wisdm_transformed.csv.gz  byte-identical gzip of Main/wisdm_main_ver_0.0/data/wisdm_data.csv — the DATA
                            file the reference reads at Main/main.py:18-20 (input fixture of configs 1-2)
  golden.json               numbers copied from the reference's own Spark output
                            (Main/wisdm_main_ver_0.0/main_result/result.txt, additional_param.csv), each
                            with the file:line it was read from
"""
import gzip
import json
import os
import re

REF = "/root/reference/Main/wisdm_main_ver_0.0"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    raw = open(os.path.join(REF, "data", "wisdm_data.csv"), "rb").read()
    with open(os.path.join(HERE, "wisdm_transformed.csv.gz"), "wb") as fh:
        with gzip.GzipFile(fileobj=fh, mode="wb", mtime=0, compresslevel=9) as gz:
            gz.write(raw)
    res = open(os.path.join(REF, "main_result", "result.txt")).read().splitlines()
    add = open(os.path.join(REF, "main_result", "additional_param.csv")).read().splitlines()

    def line(no):                      # 1-based, as cited
        return res[no - 1]

    def num(no):
        return float(re.findall(r"[-+]?\d*\.?\d+(?:[eE][-+]?\d+)?", line(no).split(":")[-1].split("=")[-1])[-1])

    dt_row = add[2].split(",")
    g = {
        "schema_lines": {"result.txt:4-18": res[3:18]},
        "activity_counts": {"result.txt:36-41": [l.strip("|").split("|") for l in res[35:41]]},
        "train_count": {"result.txt:105": int(line(105).split(":")[1])},
        "test_count": {"result.txt:106": int(line(106).split(":")[1])},
        "train_head": {"result.txt:110-114": res[109:114]},
        "test_head": {"result.txt:121-125": res[120:125]},
        "dt_model_str": {"result.txt:231": line(231)},
        "dt_top5": {"result.txt:237-241": res[236:241]},
        "dt_binary_raw": {"result.txt:248": num(248)},
        "dt_area_under_pr": {"result.txt:249": num(249)},
        "dt_area_under_roc": {"result.txt:250": num(250)},
        "dt_f1": {"result.txt:254": num(254)},
        "dt_weighted_precision": {"result.txt:255": num(255)},
        "dt_weighted_recall": {"result.txt:256": num(256)},
        "dt_accuracy": {"result.txt:257": num(257)},
        "dt_rmse": {"result.txt:261": num(261)},
        "dt_r2": {"result.txt:263": num(263)},
        "dt_mae": {"result.txt:264": num(264)},
        "dt_total": {"result.txt:268": int(num(268))},
        "dt_correct": {"result.txt:269": int(num(269))},
        "dt_wrong": {"result.txt:270": int(num(270))},
        "dt_f1_full": {"additional_param.csv:3": float(dt_row[-4])},
        "dt_accuracy_full": {"additional_param.csv:3": float(dt_row[-1])},
        "rf_model_str": {"result.txt:276": line(276)},
        "rf_correct": {"result.txt:314": int(num(314))},
        "rf_accuracy": {"result.txt:302": num(302)},
        "describe": {"result.txt:47-57": res[46:57]},
    }
    with open(os.path.join(HERE, "golden.json"), "w") as fh:
        json.dump(g, fh, indent=1)
    print("wrote fixtures:", os.path.getsize(os.path.join(HERE, "wisdm_transformed.csv.gz")), "bytes gz")


if __name__ == "__main__":
    main()
