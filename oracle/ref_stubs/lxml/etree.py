"""stdlib ElementTree dressed as the tiny part of lxml.etree the reference touches
(torch_humanoid_batch.py:50-55). Test infrastructure only."""
import xml.etree.ElementTree as _ET

Element = _ET.Element
SubElement = _ET.SubElement
ElementTree = _ET.ElementTree


class XMLParser:
    def __init__(self, *a, **k):
        pass


def _patch(elem):
    for e in elem.iter():
        pass
    return elem


if not hasattr(_ET.Element, "getchildren"):
    # python>=3.9 removed getchildren(); C Element type is immutable -> wrap on parse
    pass


class _Wrap:
    def __init__(self, e):
        self._e = e

    def __getattr__(self, n):
        return getattr(self._e, n)

    def find(self, p):
        r = self._e.find(p)
        return None if r is None else _Wrap(r)

    def findall(self, p):
        return [_Wrap(x) for x in self._e.findall(p)]

    def getchildren(self):
        return [_Wrap(x) for x in list(self._e)]


class _Tree:
    def __init__(self, t):
        self._t = t

    def getroot(self):
        return _Wrap(self._t.getroot())


def parse(src, parser=None):
    return _Tree(_ET.parse(src))
