# One image for master, agent and tasks (reference: master/Dockerfile, agent/Dockerfile, the task environment images).
#   docker build -t determined-b200 .
#   docker run -p 8080:8080 determined-b200 determined-master --port 8080
#   docker run --gpus all determined-b200 determined-agent --master-url http://master:8080
FROM nvidia/cuda:12.9.0-devel-ubuntu24.04
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip python3-venv g++ make && rm -rf /var/lib/apt/lists/*
RUN python3 -m venv /opt/det && /opt/det/bin/pip install --no-cache-dir torch numpy pyyaml requests tabulate
ENV PATH=/opt/det/bin:$PATH
WORKDIR /opt/determined-b200
COPY . .
RUN make build PY=python && pip install --no-cache-dir --no-deps .
EXPOSE 8080
CMD ["determined-master", "--port", "8080"]
