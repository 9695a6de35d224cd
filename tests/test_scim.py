"""SCIM 2.0 provisioning and the OAuth client-credentials grant, the way an identity provider drives them."""
import pytest
import requests

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import config as mconfig
from determined_b200.master import core

USER = "urn:ietf:params:scim:schemas:core:2.0:User"


@pytest.fixture()
def master(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    yield m
    m.stop()
    srv.stop()


def test_scim_is_off_by_default_and_needs_credentials(master):
    m = master
    assert requests.get(m.url + "/scim/v2/Users").status_code == 404
    with pytest.raises(ValueError):
        m.apply_config(mconfig.load(None, {}, {"scim.enabled": True}))
    cfg = mconfig.load(None, {}, {"scim.enabled": True, "scim.auth.username": "idp", "scim.auth.password": "s3cret"})
    assert cfg.printable()["scim"]["auth"]["password"] != "s3cret"
    m.apply_config(cfg)
    r = requests.get(m.url + "/scim/v2/Users")
    assert r.status_code == 401 and r.json()["schemas"] == ["urn:ietf:params:scim:api:messages:2.0:Error"] and r.json()["status"] == "401"
    assert requests.get(m.url + "/scim/v2/Users", auth=("idp", "wrong")).status_code == 401
    assert requests.get(m.url + "/scim/v2/ServiceProviderConfig").json()["patch"]["supported"] is True


def test_user_lifecycle_through_scim(master):
    m = master
    m.apply_config(mconfig.load(None, {}, {"scim.enabled": True, "scim.auth.username": "idp", "scim.auth.password": "s3cret"}))
    idp = requests.Session()
    idp.auth = ("idp", "s3cret")
    base = m.url + "/scim/v2"
    doc = {"schemas": [USER], "userName": "carol@example.com", "externalId": "00u1", "name": {"givenName": "Carol", "familyName": "Jones"},
           "emails": [{"value": "carol@example.com", "primary": True}], "active": True}
    r = idp.post(base + "/Users", json=doc)
    assert r.status_code == 201, r.text
    u = r.json()
    assert u["userName"] == "carol@example.com" and u["displayName"] == "Carol Jones" and u["active"] and u["externalId"] == "00u1"
    assert u["meta"]["location"].endswith("/scim/v2/Users/" + u["id"]) and len(u["id"]) == 32
    assert idp.post(base + "/Users", json=doc).status_code == 409 and idp.post(base + "/Users", json={}).status_code == 400

    # the provisioned user is a remote user: visible to admins, no password sign-in
    admin = api.authentication.login(m.url, "admin", "")
    row = admin.get("/api/v1/users/carol@example.com/by-username").json()["user"]
    assert row["remote"] is True and row["displayName"] == "Carol Jones"
    with pytest.raises(api.errors.APIException):
        api.authentication.login(m.url, "carol@example.com", "")

    found = idp.get(base + "/Users", params={"filter": 'userName eq "carol@example.com"'}).json()
    assert found["totalResults"] == 1 and found["Resources"][0]["id"] == u["id"]
    assert idp.get(base + "/Users", params={"filter": 'externalId eq "00u1"'}).json()["totalResults"] == 1
    assert idp.get(base + "/Users", params={"filter": "emails co x"}).status_code == 400
    everyone = idp.get(base + "/Users", params={"startIndex": 2, "count": 1}).json()
    assert everyone["totalResults"] == 3 and everyone["itemsPerPage"] == 1 and everyone["startIndex"] == 2
    assert idp.get(base + f"/Users/{u['id']}").json()["emails"][0]["value"] == "carol@example.com"
    assert idp.get(base + "/Users/nope").status_code == 404

    # deactivation signs the user out everywhere; PATCH in both spellings; PUT replaces
    m.db.exec("INSERT INTO sessions(token,user_id,expiry) VALUES('tok-carol',?,?)", (row["id"], 9e12))
    p = idp.patch(base + f"/Users/{u['id']}", json={"schemas": ["urn:ietf:params:scim:api:messages:2.0:PatchOp"], "Operations": [{"op": "Replace", "value": {"active": False}}]})
    assert p.status_code == 200 and p.json()["active"] is False
    assert m.db.one("SELECT token FROM sessions WHERE token='tok-carol'") is None
    p = idp.patch(base + f"/Users/{u['id']}", json={"Operations": [{"op": "replace", "path": "active", "value": "true"}, {"op": "replace", "path": "displayName", "value": "C. Jones"}]})
    assert p.json()["active"] is True and p.json()["displayName"] == "C. Jones"
    assert idp.patch(base + f"/Users/{u['id']}", json={"Operations": [{"op": "remove", "path": "emails"}]}).status_code == 400
    r = idp.put(base + f"/Users/{u['id']}", json=dict(doc, userName="carol.jones@example.com", displayName="Carol J."))
    assert r.json()["userName"] == "carol.jones@example.com" and r.json()["displayName"] == "Carol J."
    assert idp.put(base + f"/Users/{u['id']}", json={"userName": "admin"}).status_code == 409
    assert idp.delete(base + f"/Users/{u['id']}").status_code == 204
    assert admin.get("/api/v1/users/carol.jones@example.com/by-username").json()["user"]["active"] is False
    assert idp.get(base + "/Groups").json()["totalResults"] == 0


def test_oauth_client_credentials_open_scim(master):
    m = master
    m.apply_config(mconfig.load(None, {}, {"scim.enabled": True, "scim.auth.type": "oauth"}))
    admin = api.authentication.login(m.url, "admin", "")
    client = admin.post("/oauth2/clients", json={"name": "okta", "domain": "https://idp.example.com"}).json()
    bad = requests.post(m.url + "/oauth2/token", data={"grant_type": "client_credentials"}, auth=(client["id"], "wrong"))
    assert bad.status_code == 401
    assert requests.post(m.url + "/oauth2/token", data={"grant_type": "password"}, auth=(client["id"], client["secret"])).status_code == 400
    tok = requests.post(m.url + "/oauth2/token", data={"grant_type": "client_credentials"}, auth=(client["id"], client["secret"])).json()
    assert tok["token_type"] == "bearer" and tok["expires_in"] > 0
    r = requests.get(m.url + "/scim/v2/Users", headers={"Authorization": f"Bearer {tok['access_token']}"})
    assert r.status_code == 200 and r.json()["totalResults"] == 2
    assert requests.get(m.url + "/scim/v2/Users", headers={"Authorization": "Bearer nope"}).status_code == 401
    form = requests.post(m.url + "/oauth2/token", data={"grant_type": "client_credentials", "client_id": client["id"], "client_secret": client["secret"]})
    assert form.status_code == 200 and form.json()["access_token"] != tok["access_token"]
