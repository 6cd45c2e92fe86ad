"""Case definitions shared by tests/golden/make_golden.py (which runs the unmodified
reference on them) and the tests that replay the stored outputs."""
import json

import numpy as np

from torchkafka_b200.testing import synthetic as syn
from torchkafka_b200.testing.stub_kafka import RecordLog

CASES = {
    # BASELINE.json configs at sizes that keep the fixtures small
    "cfg1_readme_8floats": {"cfg": 1, "n": 203, "seed": 11, "kind": "json", "dtype": "float32",
                            "batch_size": 4, "partitions": 64},
    "cfg2_varlen_filter": {"cfg": 2, "n": 400, "seed": 12, "kind": "json", "dtype": "float32",
                           "min_size": 32, "batch_size": 64, "pad": True, "partitions": 64},
    "cfg2_varlen_filter_drop_last": {"cfg": 2, "n": 300, "seed": 13, "kind": "json", "dtype": "float32",
                                     "min_size": 32, "batch_size": 64, "pad": True, "drop_last": True,
                                     "partitions": 7, "layout": "chunked", "chunk": 23},
    "cfg3_tokens": {"cfg": 3, "n": 90, "seed": 14, "kind": "json", "dtype": "int32",
                    "batch_size": 16, "pad": True, "partitions": 64},
    "cfg4_images": {"cfg": 4, "n": 9, "seed": 15, "kind": "image", "hwc": (32, 24, 3), "batch_size": 4,
                    "partitions": 3, "gen": {"h": 32, "w": 24, "c": 3}},
    "cfg5_raw64": {"cfg": 5, "n": 1000, "seed": 16, "kind": "raw", "batch_size": 64, "partitions": 64},
    # SURVEY.md §4.1-3 example
    "survey_commit_example": {"custom": "survey", "kind": "json", "dtype": "float32", "min_size": 4,
                              "batch_size": 3, "pad": True},
}


def build_log(case) -> RecordLog:
    if case.get("custom") == "survey":
        lens = [4, 4, 2, 2, 2, 4, 2, 4, 4, 4, 2, 4, 4, 2]
        vals = [json.dumps([float(i)] * n).encode() for i, n in enumerate(lens)]
        i = np.arange(len(vals))
        return RecordLog(vals, i % 3, i // 3)
    return syn.make_log(case["cfg"], case["n"], seed=case["seed"], partitions=case["partitions"],
                        layout=case.get("layout", "roundrobin"), chunk=case.get("chunk", 500),
                        **case.get("gen", {}))


# SURVEY.md §4.1-11 numeric known-answer set (payloads are latin-1 text of the record bytes)
KNOWN_ANSWER_PAYLOADS = [
    "[0.1, 1e23, 8.41e21]", "[1.0000000596046448]", "[1152921573326323713]", "[16777217]",
    "[2.2250738585072011e-308, 4.9e-324]", "[2.4703282292062327e-324, 2.4703282292062328e-324]",
    "[1e400, 1E-400]", "[-0, -0.0, 0, 0.0]", "[NaN, Infinity, -Infinity]", "[1.5, 2.9, -2.9]",
    "[2147483647, -2147483648]", "[2147483648]", "[-2147483649]", "[2147483647.5]", "[1e30]",
    "[9223372036854775807]", "[9223372036854775808]", "[9.3e18]", "[1e5, 1E+5, 1e-5]",
    "[123456789012345678901234567890]", "[9007199254740993, 9007199254740992.5]",
    "[1.7976931348623157e308, 1.7976931348623158e308, 1.7976931348623159e308]",
    "[1.00000000000000011102230246251565404236316680908203125]",
    "[1.000000000000000111022302462515654042363166809082031251]",
    "[1.00000000000000011102230246251565404236316680908203124999]",
    "[1" + "0" * 308 + "]", "[1" + "0" * 309 + "]",
    "[]", "  [ 1 ,\t2 ,\n3 ]\r\n ", "ï»¿[1]",
    "[1,]", "[1.]", "[01]", "[+1]", "[1 2]", "[-]", "[1]x", "", "[1, 2", "[1e]", "[-NaN]", "[1,,2]",
    "[,1]", "[true, false]", "[null]", "[\"a\"]", "[[1, 2], [3, 4]]", "3.5", "{\"a\": 1}",
]
