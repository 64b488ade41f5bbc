"""Import-time stand-in for `easydict` (absent here). Test infrastructure only.

The reference checks `"dof_pos" in motions[0].__dict__` (motion_lib_base.py:365,371),
so the attribute dict must be the instance __dict__ itself.
"""


class EasyDict(dict):
    def __init__(self, d=None, **kwargs):
        super().__init__()
        self.__dict__ = self
        if d:
            for k, v in dict(d).items():
                self[k] = v
        for k, v in kwargs.items():
            self[k] = v
