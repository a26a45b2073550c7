"""`.voc` builders of Docodo.NET (Dict.cs:109-296) over the library's own stemmers (dcd_stem_host: the host build of the
device Snowball sources), SURVEY.md 8f row 4.

    VocBuilder                      SortedDictionary<stem, group> + AddWordsGroup + build()      Dict.cs:109-208
    OpenCorporaVocBuilder           <lemma> elements of an OpenCorpora XML dump, Russian stemmer  Dict.cs:211-256
    FreeLibVocBuilder               "form lemma tag" lines of the FreeLing word lists, English    Dict.cs:258-296

The bytes written are the format Vocab.Load reads (Dict.cs:71-95): {BinaryWriter string, int32 group}* in ordinal key order.
"""
import os
import struct

from .index import STEM_EN, STEM_NONE, STEM_RU, stem_host

GROUP_NOT_EXCACT_WORD_MASK = 0x01000000   # Dict.cs:15
GROUP_NUMBER_MASK = 0xFFFFFF              # Dict.cs:16


def _ordinal(s):
    """Index.IndexComparer == String.CompareOrdinal: UTF-16 code unit order."""
    return s.encode("utf-16-be", "surrogatepass")


class VocBuilder(dict):
    """Dict.cs:109-208.  Derive and call AddWordsGroup with morphologically grouped words, then build()."""

    def __init__(self, stemmer=STEM_NONE):
        super().__init__()
        self.stemmer = stemmer
        self.nG = 1
        self.replaces = {}
        self._stems = {}

    def _stem(self, word):
        if self.stemmer == STEM_NONE:
            return word
        r = self._stems.get(word)
        if r is None:
            r = self._stems[word] = stem_host(self.stemmer, word)
        return r

    def AddWordsGroup(self, grouplist):   # Dict.cs:120-190
        currNG = self.nG
        hasmatch = False
        found = False
        replaceGroups = set()
        for word in grouplist:
            stemme = self._stem(word)
            if not hasmatch and stemme in grouplist:
                hasmatch = True
            if stemme in self:
                newcurrNG = self[stemme]
                newcurrNG = self.replaces.get(newcurrNG, newcurrNG)
                if (currNG & GROUP_NUMBER_MASK) != (newcurrNG & GROUP_NUMBER_MASK):
                    if found:
                        replaceGroups.add(newcurrNG & GROUP_NUMBER_MASK)   # need to combine groups
                    else:
                        currNG = newcurrNG
                    found = True
        if (currNG & GROUP_NOT_EXCACT_WORD_MASK) == 0:
            hasmatch = True
        if hasmatch:
            currNG &= ~GROUP_NOT_EXCACT_WORD_MASK
        for gr in sorted(replaceGroups):
            if gr in self.replaces:
                if self.replaces[gr] != currNG:
                    raise Exception("Error duplicate replaces")
            else:
                self.replaces[gr] = currNG
        for word in grouplist:
            stemme = self._stem(word)
            if stemme not in self:
                self[stemme] = currNG
            elif hasmatch and (self[stemme] & GROUP_NOT_EXCACT_WORD_MASK) != 0:
                self[stemme] = currNG & ~GROUP_NOT_EXCACT_WORD_MASK
        self.nG += 1

    def tobytes(self):   # build(), Dict.cs:192-208
        out = bytearray()
        for key in sorted(self, key=_ordinal):
            b = key.encode("utf-8", "surrogatepass")
            n = len(b)
            while n >= 0x80:   # BinaryWriter.Write(string): 7-bit encoded byte length
                out.append((n & 0x7F) | 0x80)
                n >>= 7
            out.append(n)
            out += b
            v = self[key]
            out += struct.pack("<i", self.replaces.get(v, v))
        return bytes(out)

    def build(self, outfile):
        with open(outfile, "wb") as f:
            f.write(self.tobytes())


class FreeLibVocBuilder(VocBuilder):
    """Dict.cs:258-296: every line "form lemma ..." of every file of the folder is a group of two."""

    def __init__(self):
        super().__init__(STEM_EN)

    def AddLines(self, lines):
        for l in lines:
            a = l.split(" ")
            if len(a) >= 2 and len(a[0]) > 0 and len(a[1]) > 0:
                self.AddWordsGroup(a[:2])

    @staticmethod
    def CreateFromFolder(folder, outfile=None):
        voc = FreeLibVocBuilder()
        for name in sorted(os.listdir(folder)):   # Directory.GetFiles order is unspecified; sorted here
            path = os.path.join(folder, name)
            if os.path.isfile(path):
                voc.AddLines(open(path, encoding="utf-8", errors="replace").read().splitlines())
        if outfile:
            voc.build(outfile)
        return voc


class OpenCorporaVocBuilder(VocBuilder):
    """Dict.cs:211-256: the `t` attributes of everything inside a <lemma> element form one group."""

    def __init__(self):
        super().__init__(STEM_RU)

    @staticmethod
    def CreateFromOpenCorpora(file, outfile=None):
        import xml.etree.ElementTree as ET
        builder = OpenCorporaVocBuilder()
        grouplist, inside = [], False
        for ev, el in ET.iterparse(file, events=("start", "end")):
            if el.tag == "lemma":
                if ev == "start":
                    grouplist, inside = [], True
                    # XmlReader.Read visits the lemma element itself first; it has no `t` attribute in the dump format
                else:
                    builder.AddWordsGroup(grouplist)
                    inside = False
                    el.clear()
            elif inside and ev == "start":
                t = el.get("t")
                if t:
                    grouplist.append(t)
        if outfile:
            builder.build(outfile)
        return builder
