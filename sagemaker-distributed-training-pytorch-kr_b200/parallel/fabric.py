"""placeholder replaced below"""
class _Null:
    usable = False
def get_fabric(group=None):
    return _Null()
