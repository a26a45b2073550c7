"""Document sets shared by the oracle tests and the GPU parity tests.

A document set is a list of (source, doc_name, [(page_id, text), ...]) in indexing order --
the output contract of the reference's ingest layer (DataSources.cs:34-126).
"""
import os

PAGE_SIZE = 3000  # DataSources.cs:308


def indextest_source(golden, npages):
    """TestDataSource of XUnitDocodoTest/IndexTest.cs:87-124."""
    docs = []
    for name in ("Sample", "Dump"):
        pages = [("0", golden["TestHeaders1"] + "Name=%s\n" % name), ("1", golden["TestText1"])]
        pages += [(str(q + 2), golden["TestText1"]) for q in range(npages - 1)]
        docs.append(("Test", name, pages))
    return docs


def text_file_source(path, encoding="cp1251", source="doc", name=None):
    """IndexTextFilesDataSource / IndexedTextFile (DataSources.cs:305-459): header page "0" then
    3000-char pages "1".. ; the charset is fixed explicitly."""
    name = name or os.path.basename(path)
    text = open(path, "rb").read().decode(encoding)
    pages = [("0", "Name=%s\nSource=%s\n" % (name, source))]
    for k in range(0, len(text), PAGE_SIZE):
        pages.append((str(k // PAGE_SIZE + 1), text[k:k + PAGE_SIZE]))
    return [(source, name, pages)]


def feed(index, docs):
    """Run IndexTask's loop (Build.cs:460-551) over a document set on any Index-like object."""
    for source, name, pages in docs:
        index.add_doc(source, name)
        for pid, text in pages:
            index.index_page(pid, text)
