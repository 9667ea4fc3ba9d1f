class TensorLayout:
    """reference quest/utils/utils.py:1-5"""
    NHD = 0
    HND = 1

    FORMAT2STR = {0: "NHD", 1: "HND"}
