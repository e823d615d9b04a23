class Poly3DCollection:
    def __init__(self, *a, **k):
        pass
