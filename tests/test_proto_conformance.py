"""The API definition (``determined_b200/proto``), the generated bindings and the master agree.

* the proto reader: grammar subset, name resolution, errors;
* the generated module: enum member naming, construction / ``from_json`` / ``to_json`` semantics the reference's callers
  rely on (unset vs None, required keywords), freshness of the committed file;
* every rpc is served by a route of the master;
* a live cluster is driven THROUGH the typed functions, every answer the master gives (to this test and to the trial
  processes it starts) is recorded in the dispatcher and parsed strictly with the message its rpc declares.
(reference counterpart: the swagger-generated ``bindings.py`` is exercised by every e2e test; here one test owns the contract.)
"""
import base64
import collections
import pathlib
import re
import threading
import time

import pytest
import yaml

from determined_b200.common import api
from determined_b200.common.api import bindings as b
from determined_b200.common.api import generate_bindings, protolite
from determined_b200.master import api as mapi
from determined_b200.master import core, openapi

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"


# --------------------------------------------------------------------------------------------------------------------- reader
SMALL = '''
syntax = "proto3";
package acme.zoo.v1;
import "google/protobuf/struct.proto";
// Kinds of animal.
enum Kind { KIND_UNSPECIFIED = 0; KIND_CAT = 1; KIND_DOG = 2; }
// An animal.
message Animal {
  option (required) = "id, kind";
  // Id.
  int32 id = 1;
  Kind kind = 2;
  repeated string nick_names = 3;
  map<string, Animal> friends = 4;
  google.protobuf.Struct extra = 5;
  uint64 big_number = 6;
  string raw_key = 7 [json_name = "raw_key"];
  oneof home { string zoo_name = 8; int32 cage = 9; }
  message Tag { string text = 1; }
  Tag tag = 10;
}
message GetAnimalRequest { int32 animal_id = 1; bool verbose = 2; }
message GetAnimalResponse { Animal animal = 1; }
service Zoo {
  // Look one up.
  rpc GetAnimal(GetAnimalRequest) returns (GetAnimalResponse) { option (http) = { get: "/zoo/{animal_id}" }; }
  rpc PutAnimal(Animal) returns (GetAnimalResponse) { option (http) = { put: "/zoo/{id}" body: "*" }; }
}
'''


def test_reader_parses_the_subset_and_resolves_names():
    reg = protolite.Registry()
    reg.add_file(SMALL, "zoo.proto")
    reg.resolve()
    a = reg.messages["acme.zoo.v1.Animal"]
    by = {f.name: f for f in a.fields}
    assert a.doc == "An animal." and a.required == ["id", "kind"] and by["id"].doc == "Id."
    assert by["kind"].type == "acme.zoo.v1.Kind" and by["nick_names"].repeated and by["friends"].map_value
    assert by["friends"].type == "acme.zoo.v1.Animal" and by["tag"].type == "acme.zoo.v1.Animal.Tag"
    assert by["nick_names"].json == "nickNames" and by["raw_key"].json == "raw_key"
    assert by["zoo_name"].oneof == "home" and by["cage"].oneof == "home"
    assert reg.binding_name("acme.zoo.v1.Animal.Tag") == "v1AnimalTag" and reg.binding_name("acme.zoo.v1.Kind") == "v1Kind"
    get, put = reg.rpcs
    assert (get.method, get.path, get.body, get.doc) == ("get", "/zoo/{animal_id}", None, "Look one up.")
    assert (put.method, put.body, put.request) == ("put", "*", "acme.zoo.v1.Animal")
    sch = reg.schemas()
    assert sch["v1Animal"]["required"] == ["id", "kind"]
    assert sch["v1Animal"]["properties"]["bigNumber"] == {"type": "string", "format": "uint64"}
    assert sch["v1Animal"]["properties"]["friends"]["additionalProperties"] == {"$ref": "#/components/schemas/v1Animal"}
    assert sch["v1Kind"]["enum"] == ["KIND_UNSPECIFIED", "KIND_CAT", "KIND_DOG"]


@pytest.mark.parametrize("text,needle", [
    ('package a.v1; message M { Missing x = 1; }', "not defined"),
    ('package a.v1; message M { option (required) = "nope"; int32 x = 1; }', "does not exist"),
    ('package a.v1; message R {} message S {} service X { rpc Y(R) returns (S) { option (http) = { get: "/y/{zzz}" }; } }', "path parameter"),
    ('package a.v1; message R {} message S {} service X { rpc Y(R) returns (S); }', "no (http) option"),
    ('package a.v1; message M { int32 x = ; }', "field number"),
])
def test_reader_reports_mistakes(text, needle):
    reg = protolite.Registry()
    with pytest.raises(protolite.ProtoError, match=re.escape(needle)):
        reg.add_file(text, "bad.proto")
        reg.resolve()


def test_two_packages_with_the_same_local_name_get_their_package_in_front():
    reg = protolite.load()
    assert reg.binding_name("determined.experiment.v1.State") == "experimentv1State"
    assert reg.binding_name("determined.checkpoint.v1.State") == "checkpointv1State"
    assert reg.binding_name("determined.experiment.v1.TrialExitedEarly.ExitedReason") == "v1TrialExitedEarlyExitedReason"


# ------------------------------------------------------------------------------------------------------------------- bindings
def test_committed_bindings_are_fresh():
    assert generate_bindings.main(["--check"]) == 0


def test_enum_members_drop_the_common_prefix_and_travel_by_wire_name():
    assert b.experimentv1State.COMPLETED.value == "STATE_COMPLETED"
    assert b.v1TrialExitedEarlyExitedReason.INVALID_HP.value == "EXITED_REASON_INVALID_HP"
    assert b.experimentv1State("STATE_PAUSED") is b.experimentv1State.PAUSED
    assert b.v1LogLevel.WARNING.to_json() == "LOG_LEVEL_WARNING"


def test_message_semantics():
    with pytest.raises(TypeError, match="requestId"):
        b.v1TrialCreated()
    with pytest.raises(TypeError, match="unexpected"):
        b.v1TrialCreated(requestId="r", nope=1)
    ev = b.v1SearcherEvent(id=3, trialCreated=b.v1TrialCreated(requestId="r"), trialClosed=None)
    assert ev.trialProgress is None                       # unset reads as None
    assert ev.to_json(omit_unset=True) == {"id": 3, "trialCreated": {"requestId": "r"}, "trialClosed": None}   # explicit None is sent
    assert set(ev.to_json()) == {"id", "initialOperations", "trialCreated", "validationCompleted", "trialClosed", "trialExitedEarly",
                                 "trialProgress", "experimentInactive"}
    back = b.v1SearcherEvent.from_json({"id": 3, "trialCreated": {"requestId": "r"}, "trialClosed": None, "futureField": 1})
    assert back == ev and back != b.v1SearcherEvent(id=4)
    with pytest.raises(KeyError, match="futureField"):
        b.v1SearcherEvent.from_json({"id": 3, "futureField": 1}, strict=True)
    with pytest.raises(KeyError, match="required"):
        b.v1SearcherEvent.from_json({})
    assert b.v1SearcherEvent.from_json({}, partial=True).id is None
    inact = b.v1SearcherEvent.from_json({"id": 9, "experimentInactive": {"experimentState": "STATE_CANCELED"}})
    assert inact.experimentInactive.experimentState is b.experimentv1State.CANCELED
    assert inact.to_json(True) == {"id": 9, "experimentInactive": {"experimentState": "STATE_CANCELED"}}
    dp = b.v1DataPoint(batches=1, value=float("inf"))
    assert dp.to_json(True) == {"batches": 1, "value": "Infinity"} and b.v1DataPoint.from_json(dp.to_json(True)).value == float("inf")
    assert "requestId='r'" in repr(ev)


def test_every_rpc_is_served_by_a_route():
    doc = openapi.build(mapi.ROUTES)
    have = {(m.upper(), openapi._shape(p)) for p, ops in doc["paths"].items() for m in ops}
    missing = [r.name for r in protolite.load().rpcs if (r.method.upper(), openapi._shape(r.path)) not in have]
    assert not missing
    op = doc["paths"]["/api/v1/experiments/{eid}"]["get"]
    assert op["x-rpc"] == "GetExperiment" and op["responses"]["200"]["content"]["application/json"]["schema"] == {"$ref": "#/components/schemas/v1GetExperimentResponse"}
    q = {p["name"] for p in doc["paths"]["/api/v1/experiments"]["get"]["parameters"] if p["in"] == "query"}
    assert {"offset", "limit", "archived", "orderBy", "projectId"} <= q
    assert doc["components"]["schemas"]["v1Experiment"]["properties"]["state"] == {"allOf": [{"$ref": "#/components/schemas/experimentv1State"}], "description": "Current state."}


# ----------------------------------------------------------------------------------------------------------------- live master
@pytest.fixture()
def recorded_cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=2, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}}, scheduler_type="priority")
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    records = []
    lock = threading.Lock()

    def rec(method, pattern, body, out):
        with lock:
            records.append((method, re.sub(r"\(\?P<(\w+)>\[\^/\]\+\)", r"{\1}", pattern), body, out))
    mapi.RECORDER = rec
    sess = api.authentication.login(m.url, "admin", "")
    try:
        yield m, sess, records
    finally:
        mapi.RECORDER = None
        m.stop()
        srv.stop()


def _files():
    return [b.v1File(path=p.name, type=48, content=base64.b64encode(p.read_bytes()).decode(), mode=0o644, mtime=int(time.time()))
            for p in FIX.iterdir() if p.is_file()]


CFG = {"name": "typed", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "single", "metric": "val_loss", "max_length": {"batches": 16}},
       "hyperparameters": {"learning_rate": 0.1, "batch_size": 8, "global_batch_size": 8}, "scheduling_unit": 8,
       "min_validation_period": {"batches": 8}, "resources": {"slots_per_trial": 1}}


def test_the_master_answers_what_the_definition_says(recorded_cluster):
    m, sess, records = recorded_cluster
    # ---- a tour through the typed functions ------------------------------------------------------------------------------
    assert b.get_GetMaster(sess).clusterName and b.get_GetMe(sess).user.username == "admin"
    assert b.get_GetMasterConfig(sess).config["resource_manager"]
    u = b.post_PostUser(sess, body=b.v1PostUserRequest(user=b.v1User(username="typed-user", admin=False, active=True), password="pw")).user
    assert b.patch_PatchUser(sess, userId=u.id, body=b.v1PatchUserRequest(userId=u.id, user=b.v1PatchUser(displayName="Typed"))).user.displayName == "Typed"
    assert any(x.username == "typed-user" for x in b.get_GetUsers(sess, limit=50).users)
    g = b.post_CreateGroup(sess, body=b.v1CreateGroupRequest(name="typed-group", addUsers=[u.id])).group
    assert b.get_GetGroup(sess, groupId=g.groupId).group.users[0].username == "typed-user"
    assert b.post_GetGroups(sess, body=b.v1GetGroupsRequest(userId=u.id)).groups[0].numMembers == 1
    role = b.post_ListRoles(sess, body=b.v1ListRolesRequest()).roles[0]
    b.post_AssignRoles(sess, body=b.v1AssignRolesRequest(userRoleAssignments=[b.v1UserRoleAssignment(userId=u.id, roleAssignment=b.v1RoleAssignment(role=role))]))
    assert b.get_GetRolesAssignedToUser(sess, userId=u.id).roles[0].role.roleId == role.roleId
    b.put_UpdateGroup(sess, groupId=g.groupId, body=b.v1UpdateGroupRequest(groupId=g.groupId, removeUsers=[u.id]))
    b.delete_DeleteGroup(sess, groupId=g.groupId)

    ws = b.post_PostWorkspace(sess, body=b.v1PostWorkspaceRequest(name="typed-ws")).workspace
    pr = b.post_PostProject(sess, workspaceId=ws.id, body=b.v1PostProjectRequest(name="typed-proj", workspaceId=ws.id, description="d")).project
    assert b.get_GetWorkspaceProjects(sess, id=ws.id).projects[0].id == pr.id and b.get_GetProject(sess, id=pr.id).project.name == "typed-proj"
    assert any(w.id == ws.id for w in b.get_GetWorkspaces(sess).workspaces) and b.get_GetWorkspace(sess, id=ws.id).workspace.name == "typed-ws"
    b.put_PutTemplate(sess, templateName="typed-tpl", body=b.v1Template(name="typed-tpl", config={"max_restarts": 1}))
    assert b.get_GetTemplate(sess, templateName="typed-tpl").template.config == {"max_restarts": 1}
    assert [t.name for t in b.get_GetTemplates(sess).templates] == ["typed-tpl"]
    hook = b.post_PostWebhook(sess, body=b.v1Webhook(url="http://127.0.0.1:9/x", webhookType=b.v1WebhookType.DEFAULT, triggers=[
        b.v1Trigger(triggerType=b.v1TriggerType.EXPERIMENT_STATE_CHANGE, condition={"state": ["COMPLETED"]})])).webhook
    assert b.get_GetWebhooks(sess).webhooks[0].triggers[0].triggerType is b.v1TriggerType.EXPERIMENT_STATE_CHANGE
    b.delete_DeleteWebhook(sess, id=hook.id)

    sim = b.post_PreviewHPSearch(sess, body=b.v1PreviewHPSearchRequest(config=yaml.safe_dump(CFG), seed=1)).simulation
    assert sim.trials[0].operations[-1].validateAfter.length in (16, "16")
    created = b.post_CreateExperiment(sess, body=b.v1CreateExperimentRequest(config=yaml.safe_dump(CFG), modelDefinition=_files(), activate=True, projectId=pr.id))
    eid = created.experiment.id
    assert created.config["searcher"]["name"] == "single"
    b.patch_PatchExperiment(sess, experimentId=eid, body=b.v1PatchExperiment(id=eid, description="patched", labels=["a"],
                                                                             resources=b.v1PatchExperimentPatchResources(priority=40)))
    deadline = time.time() + 240
    while b.get_GetExperiment(sess, experimentId=eid).experiment.state not in (b.experimentv1State.COMPLETED, b.experimentv1State.ERROR):
        assert time.time() < deadline
        b.get_GetJobs(sess)
        b.get_GetTasks(sess)
        time.sleep(0.5)
    exp = b.get_GetExperiment(sess, experimentId=eid).experiment
    assert exp.state is b.experimentv1State.COMPLETED and exp.description == "patched" and exp.labels == ["a"] and exp.projectId == pr.id
    assert [e.id for e in b.get_GetExperiments(sess, projectId=pr.id, orderBy="ORDER_BY_DESC").experiments] == [eid]
    trial = b.get_GetExperimentTrials(sess, experimentId=eid).trials[0]
    assert trial.totalBatchesProcessed == 16 and trial.state is b.experimentv1State.COMPLETED
    assert b.get_GetTrial(sess, trialId=trial.id).trial.latestValidation.totalBatches == 16
    reports = b.get_GetTrialMetrics(sess, trialId=trial.id).metrics
    assert [r.totalBatches for r in reports if r.group == "validation"] == [8, 16]
    assert b.get_GetMetrics(sess, trialIds=[trial.id], group="training").metrics
    names = b.get_ExpMetricNames(sess, ids=[eid]).results[0]
    assert "val_loss" in names.validationMetrics and b.get_GetExperimentMetricNames(sess, experimentId=eid).searcherMetric == "val_loss"
    assert b.get_TrialsSnapshot(sess, experimentId=eid, metricName="val_loss", metricType=b.v1MetricType.VALIDATION, batchesProcessed=16).trials[0].trialId == trial.id
    assert len(b.get_TrialsSample(sess, experimentId=eid, metricName="val_loss", metricType=b.v1MetricType.VALIDATION).trials[0].data) == 2
    summ = b.get_SummarizeTrial(sess, trialId=trial.id, maxDatapoints=10)
    assert {s.name for s in summ.metrics} >= {"loss", "val_loss"}
    assert b.get_CompareTrials(sess, trialIds=[trial.id]).trials[0].trial.id == trial.id
    assert b.get_GetBestSearcherValidationMetric(sess, experimentId=eid).metric is not None
    assert base64.b64decode(b.get_GetModelDef(sess, experimentId=eid).b64Tgz)
    logs = b.get_TrialLogs(sess, trialId=trial.id, limit=1000).logs
    assert logs and b.get_TaskLogs(sess, taskId=trial.taskId, limit=5).logs[0].taskId == trial.taskId
    cks = b.get_GetExperimentCheckpoints(sess, id=eid, sortBy="SORT_BY_BATCH_NUMBER", orderBy="ORDER_BY_DESC").checkpoints
    assert cks and cks[0].state is b.checkpointv1State.COMPLETED and cks[0].training.trialId == trial.id
    ck = b.get_GetCheckpoint(sess, checkpointUuid=cks[0].uuid).checkpoint
    assert ck.metadata["steps_completed"] == 16 and "state_dict.pth" in ck.resources
    assert b.get_GetTrialCheckpoints(sess, id=trial.id).pagination.total == len(cks)
    noted = b.post_PostCheckpointMetadata(sess, checkpointUuid=ck.uuid, body=b.v1PostCheckpointMetadataRequest(
        checkpoint=b.v1Checkpoint(uuid=ck.uuid, metadata={"note": "typed"}))).checkpoint
    assert noted.metadata["note"] == "typed" and noted.metadata["steps_completed"] == 16

    model = b.post_PostModel(sess, body=b.v1PostModelRequest(name="typed-model", description="m", labels=["x"], metadata={"k": 1})).model
    ver = b.post_PostModelVersion(sess, modelName="typed-model", body=b.v1PostModelVersionRequest(modelName="typed-model", checkpointUuid=ck.uuid, comment="c")).modelVersion
    assert ver.version == 1 and ver.checkpoint.uuid == ck.uuid and ver.model.id == model.id
    assert b.get_GetModelVersions(sess, modelName="typed-model").modelVersions[0].id == ver.id
    assert b.get_GetModelVersion(sess, modelName="typed-model", modelVersionNum=1).modelVersion.comment == "c"
    assert b.patch_PatchModel(sess, modelName="typed-model", body=b.v1PatchModel(description="new")).model.description == "new"
    assert [x.name for x in b.get_GetModels(sess, name="typed").models] == ["typed-model"] and b.get_GetModel(sess, modelName="typed-model").model.numVersions == 1
    b.post_ArchiveModel(sess, modelName="typed-model")
    b.post_UnarchiveModel(sess, modelName="typed-model")
    b.delete_DeleteModelVersion(sess, modelName="typed-model", modelVersionNum=1)
    b.delete_DeleteModel(sess, modelName="typed-model")

    agents = b.get_GetAgents(sess).agents
    assert agents and agents[0].slots["0"].device.type is b.v1DeviceType.CPU
    assert b.get_GetAgent(sess, agentId=agents[0].id).agent.id == agents[0].id and len(b.get_GetSlots(sess, agentId=agents[0].id).slots) == 2
    assert not b.post_DisableSlot(sess, agentId=agents[0].id, slotId="1", body=b.v1DisableSlotRequest()).agent.slots["1"].enabled
    assert b.post_EnableSlot(sess, agentId=agents[0].id, slotId="1").agent.slots["1"].enabled
    pool = b.get_GetResourcePools(sess).resourcePools[0]
    assert pool.schedulerType is b.v1SchedulerType.PRIORITY and pool.slotsAvailable == 2

    cmd = b.post_LaunchCommand(sess, body=b.v1LaunchCommandRequest(config={"entrypoint": ["sleep", "30"], "resources": {"slots": 0}})).command
    assert b.get_GetCommand(sess, commandId=cmd.id).command.id == cmd.id and any(c.id == cmd.id for c in b.get_GetCommands(sess).commands)
    assert b.post_KillCommand(sess, commandId=cmd.id).command.id == cmd.id

    b.post_ArchiveExperiment(sess, id=eid)
    b.post_UnarchiveExperiment(sess, id=eid)

    # ---- the wider surface (master/api_rest.py) --------------------------------------------------------------------------------
    assert b.post_ArchiveExperiments(sess, body=b.v1ArchiveExperimentsRequest(experimentIds=[eid, 424242])).results[1].error
    assert b.post_UnarchiveExperiments(sess, body=b.v1UnarchiveExperimentsRequest(filters=b.v1BulkExperimentFilters(name="typed", projectId=pr.id))).results[0].id == eid
    for fn, req in ((b.post_ActivateExperiments, b.v1ActivateExperimentsRequest), (b.post_PauseExperiments, b.v1PauseExperimentsRequest),
                    (b.post_CancelExperiments, b.v1CancelExperimentsRequest), (b.post_KillExperiments, b.v1KillExperimentsRequest)):
        assert fn(sess, body=req(experimentIds=[eid])).results[0].id == eid
    assert b.post_MoveExperiments(sess, body=b.v1MoveExperimentsRequest(experimentIds=[eid], destinationProjectId=pr.id)).results[0].error == ""
    assert b.put_PutExperimentLabel(sess, experimentId=eid, label="typed-label").labels == ["a", "typed-label"]
    assert "typed-label" in b.get_GetExperimentLabels(sess, projectId=pr.id).labels
    assert b.delete_DeleteExperimentLabel(sess, experimentId=eid, label="typed-label").labels == ["a"]
    tree = b.get_GetModelDefTree(sess, experimentId=eid).files
    assert any(f.name == "model_def.py" and not f.isDir for f in tree)
    assert base64.b64decode(b.post_GetModelDefFile(sess, experimentId=eid, body=b.v1GetModelDefFileRequest(experimentId=eid, path="model_def.py")).file)
    hist = b.get_GetExperimentValidationHistory(sess, experimentId=eid).validationHistory
    assert hist and hist[0].trialId == trial.id
    assert b.get_CurrentUser(sess).user.username == "admin" and b.get_GetUser(sess, userId=u.id).user.username == "typed-user"
    assert b.get_GetUserByUsername(sess, username="typed-user").user.id == u.id
    assert b.patch_PatchUsers(sess, body=b.v1PatchUsersRequest(userIds=[u.id], activate=True)).results[0].error == ""
    g2 = b.post_CreateGroup(sess, body=b.v1CreateGroupRequest(name="typed-group-2")).group
    b.patch_AssignMultipleGroups(sess, body=b.v1AssignMultipleGroupsRequest(userIds=[u.id], addGroups=[g2.groupId], removeGroups=[]))
    assert b.get_GetGroup(sess, groupId=g2.groupId).group.users[0].id == u.id
    b.post_PostUserSetting(sess, body=b.v1PostUserSettingRequest(settings=[b.v1UserWebSetting(key="k", storagePath="p", value="v")]))
    assert b.get_GetUserSetting(sess).settings[0].value == "v"
    b.post_ResetUserSetting(sess)
    assert b.post_AddProjectNote(sess, projectId=pr.id, body=b.v1Note(name="n", contents="c")).notes[0].name == "n"
    assert b.put_PutProjectNotes(sess, projectId=pr.id, body=b.v1PutProjectNotesRequest(projectId=pr.id, notes=[])).notes == []
    assert any(c.column == "hp.learning_rate" for c in b.get_GetProjectColumns(sess, id=pr.id).columns)
    assert any(r.metricsName == "validation.val_loss" for r in b.get_GetProjectNumericMetricsRange(sess, id=pr.id).ranges)
    ws2 = b.post_PostWorkspace(sess, body=b.v1PostWorkspaceRequest(name="typed-ws-2")).workspace
    pr2 = b.post_PostProject(sess, workspaceId=ws.id, body=b.v1PostProjectRequest(name="typed-proj-2", workspaceId=ws.id)).project
    b.post_MoveProject(sess, projectId=pr2.id, body=b.v1MoveProjectRequest(projectId=pr2.id, destinationWorkspaceId=ws2.id))
    assert b.get_GetProject(sess, id=pr2.id).project.workspaceId == ws2.id
    b.post_PinWorkspace(sess, id=ws2.id)
    b.post_UnpinWorkspace(sess, id=ws2.id)
    b.post_PostModel(sess, body=b.v1PostModelRequest(name="typed-model-2", labels=["lab"]))
    b.post_MoveModel(sess, modelName="typed-model-2", body=b.v1MoveModelRequest(modelName="typed-model-2", destinationWorkspaceId=ws2.id))
    assert b.get_GetModel(sess, modelName="typed-model-2").model.workspaceId == ws2.id and b.get_GetModelLabels(sess, workspaceId=ws2.id).labels == ["lab"]
    assert b.get_GetSlot(sess, agentId=agents[0].id, slotId="0").slot.id == "0"
    assert b.get_GetTelemetry(sess).enabled is False and b.get_GetActiveTasksCount(sess).commands >= 0
    assert b.get_GetTask(sess, taskId=trial.taskId).task.allocations and b.get_GetTask(sess, taskId=cmd.id).task.taskType == "TASK_TYPE_COMMAND"
    assert b.get_GetJobQueueStats(sess).results[0].resourcePool == "default"
    assert b.delete_DeleteExperiments(sess, experimentIds=[424242]).results[0].error
    b.delete_DeleteCheckpoints(sess, body=b.v1DeleteCheckpointsRequest(checkpointUuids=[cks[-1].uuid]))
    b.delete_DeleteTemplate(sess, templateName="typed-tpl")
    with pytest.raises(api.errors.NotFoundException):
        b.get_GetExperiment(sess, experimentId=987654)
    b.post_Logout(sess)

    # ---- everything the master said, against the definition -----------------------------------------------------------------
    mapi.RECORDER = None
    reg = protolite.load()
    rpcs = {(r.method.upper(), openapi._shape(r.path)): r for r in reg.rpcs}
    seen = collections.Counter()
    problems = []
    for method, path, body, out in list(records):
        r = rpcs.get((method, openapi._shape(path)))
        if r is None or isinstance(out, tuple):
            continue
        seen[r.name] += 1
        try:
            getattr(b, reg.binding_name(r.response)).from_json(out, strict=True)
        except Exception as e:  # noqa
            problems.append(f"{r.name} response: {type(e).__name__}: {e}")
        if r.body and isinstance(body, dict):
            try:
                if r.body == "*":
                    getattr(b, reg.binding_name(r.request)).from_json(body, strict=True, partial=True)
                else:
                    f = next(x for x in reg.messages[r.request].fields if x.name == r.body)
                    if f.type in reg.messages:
                        getattr(b, reg.binding_name(f.type)).from_json(body, strict=True, partial=True)
            except Exception as e:  # noqa
                problems.append(f"{r.name} request: {type(e).__name__}: {e}")
    assert not problems, "\n".join(sorted(set(problems)))
    # the harness inside the trial process talked to these
    for name in ("GetCurrentTrialSearcherOperation", "CompleteTrialSearcherValidation", "ReportTrialMetrics",
                 "ReportCheckpoint", "PostTrialRunnerMetadata", "AllocationPreemptionSignal"):
        assert seen[name] > 0, f"no {name} call was recorded"
    assert len(seen) >= 115, sorted(seen)
