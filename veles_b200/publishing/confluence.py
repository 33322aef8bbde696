"""Minimal Confluence XML-RPC client for the publishing backend.

Reference: ``veles/publishing/confluence.py:46-300`` (a wrapper over ``xmlrpclib.ServerProxy(server + "/rpc/xmlrpc")``
speaking the ``confluence2`` remote API: login token, pages, attachments, labels).  Only the standard library is used;
every call goes through ``_call`` so that an expired token is renewed once.
"""
from __future__ import annotations

import mimetypes
import os
import xmlrpc.client


class ConfluenceError(RuntimeError):
    pass


class Confluence(object):
    def __init__(self, server, username, password, timeout=120, api="confluence2"):
        self.server_url = server.rstrip("/")
        self.username = username
        self.password = password
        self.api = api
        self.timeout = timeout
        self._proxy = xmlrpc.client.ServerProxy(self.server_url + "/rpc/xmlrpc", allow_none=True)
        self._token = None
        self.login()

    # -- plumbing ----------------------------------------------------------------------------------------------
    def login(self):
        try:
            self._token = getattr(self._proxy, self.api).login(self.username, self.password)
        except (xmlrpc.client.Error, OSError) as e:
            raise ConfluenceError("Confluence login at %s failed: %s" % (self.server_url, e))
        return self._token

    @property
    def token(self):
        return self._token

    def _call(self, method, *args):
        fn = getattr(getattr(self._proxy, self.api), method)
        try:
            return fn(self._token, *args)
        except xmlrpc.client.Fault as fault:
            if "token" in str(fault.faultString).lower() or "session" in str(fault.faultString).lower():
                self.login()
                return fn(self._token, *args)
            raise

    # -- pages -------------------------------------------------------------------------------------------------
    def get_page(self, page, space):
        """The page dict; raises ``xmlrpc.client.Fault`` when it does not exist."""
        return self._call("getPage", space, page)

    def get_page_summary(self, page, space):
        try:
            return self._call("getPageSummary", space, page)
        except xmlrpc.client.Fault:
            return None

    def get_page_id(self, page, space):
        summary = self.get_page_summary(page, space)
        return None if summary is None else summary["id"]

    def get_spaces(self):
        return self._call("getSpaces")

    def get_pages(self, space):
        return self._call("getPages", space)

    def convert_wiki_to_storage_format(self, markup):
        return self._call("convertWikiToStorageFormat", markup)

    def store_page_content(self, page, space, content, convert_wiki=True, parent=None):
        """Create the page or replace its body; returns the stored page dict (``id``, ``url``, ...)."""
        if convert_wiki:
            content = self.convert_wiki_to_storage_format(content)
        try:
            data = self.get_page(page, space)
        except xmlrpc.client.Fault:
            data = {"title": page, "space": space}
            if parent is not None:
                parent_id = self.get_page_id(parent, space)
                if parent_id is None:
                    raise ConfluenceError("parent page %r does not exist in space %r" % (parent, space))
                data["parentId"] = parent_id
        data["content"] = content
        return self._call("storePage", data)

    def add_label_by_name(self, label_name, object_id):
        return self._call("addLabelByName", label_name, object_id)

    # -- attachments -------------------------------------------------------------------------------------------
    def attach_file(self, page, space, files, comments=None):
        """``files``: {attachment name: path or bytes}.  Existing attachments of the same name are replaced."""
        page_id = self.get_page_id(page, space)
        if page_id is None:
            raise ConfluenceError("page %r does not exist in space %r" % (page, space))
        stored = []
        for name, src in files.items():
            if isinstance(src, (bytes, bytearray)):
                data = bytes(src)
            else:
                with open(src, "rb") as f:
                    data = f.read()
            try:
                self._call("removeAttachment", page_id, name)
            except xmlrpc.client.Fault:
                pass
            meta = {"fileName": name,
                    "contentType": mimetypes.guess_type(name)[0] or "application/binary",
                    "comment": (comments or {}).get(name, os.path.basename(name))}
            stored.append(self._call("addAttachment", page_id, meta, xmlrpc.client.Binary(data)))
        return stored

    def remove_all_attachments(self, page, space):
        page_id = self.get_page_id(page, space)
        if page_id is None:
            return 0
        n = 0
        for att in self._call("getAttachments", page_id):
            self._call("removeAttachment", page_id, att["fileName"])
            n += 1
        return n
